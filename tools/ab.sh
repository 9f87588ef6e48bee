#!/bin/bash
# A/B timing of two builds of the library on the same box: ab/libdsdf_old.so against the in-tree build, alternating.
for i in 1 2 3; do
  echo -n "old: "; DSDF_B200_LIB=$PWD/ab/libdsdf_old.so timeout 120 python tools/kernel_times.py 20 2>&1 | tail -1
  echo -n "new: "; timeout 120 python tools/kernel_times.py 20 2>&1 | tail -1
done
