#!/bin/bash
# same-box A/B of library builds: every ab_libs/*.so in turn (twice, interleaved) under profiles/profile_wide.py
set -u
L=reinforcement_learning_cpp_parallel_b200/libb200ppo.so
cp $L /tmp/lib_keep.so
for rep in 1 2; do for v in ab_libs/*.so; do
  cp $v $L; printf "%-22s " $(basename $v); timeout 120 python profiles/profile_wide.py 2>&1 | grep update_resident | tail -1 | cut -c1-60
done; done
cp /tmp/lib_keep.so $L
