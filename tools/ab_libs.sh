#!/bin/bash
# same-box A/B of library builds: every ab_libs/*.so in turn (twice, interleaved) under profiles/profile_wide.py;
# prints the timed repetitions of each run (back-to-back updates: the later ones show the sustained clock)
set -u
L=reinforcement_learning_cpp_parallel_b200/libb200ppo.so
cp $L /tmp/lib_keep.so
for rep in 1 2; do for v in ab_libs/*.so; do
  cp $v $L; printf "%-22s " $(basename $v); timeout 120 python profiles/profile_wide.py 524288 32 10 4 2>&1 | grep update_resident | cut -c17-24 | tr "\n" " "; echo
done; done
cp /tmp/lib_keep.so $L
