#!/bin/bash
# A/B timing of two builds of the library on the same box: ab/libdsdf_old.so against the in-tree build, alternating.
# Build the baseline first:  git archive <rev> deepsdf_b200/csrc include | tar -x -C /tmp/old && make -C /tmp/old/deepsdf_b200/csrc
#                            && cp /tmp/old/deepsdf_b200/libdsdf_b200.so ab/libdsdf_old.so      (ab/ is git-ignored)
[ -f ab/libdsdf_old.so ] || { echo "ab/libdsdf_old.so missing (see the header of this script)"; exit 1; }
for i in 1 2 3; do
  echo -n "old: "; DSDF_B200_LIB=$PWD/ab/libdsdf_old.so timeout 120 python tools/kernel_times.py 20 2>&1 | tail -1
  echo -n "new: "; timeout 120 python tools/kernel_times.py 20 2>&1 | tail -1
done
