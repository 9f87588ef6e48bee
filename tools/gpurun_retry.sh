#!/bin/bash
# usage: tools/gpurun_retry.sh <gpus> <logfile> <timeout> <command>   - retries while the pod answers "busy" (rc 3)
g=$1; shift; log=$1; shift; to=$1; shift
for i in $(seq 1 40); do
  if [ "$g" = "1" ]; then /usr/local/graft/bin/gpurun --timeout $to -- "$@" > $log 2>&1; else /usr/local/graft/bin/gpurun --gpus $g --timeout $to -- "$@" > $log 2>&1; fi
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 60
done
exit 3
