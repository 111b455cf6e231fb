#!/bin/bash
# quick A/B pass for a change of the wide step: its parity tests, then the configs[4] update with and without the change
set -u
mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_gpu_wide.py tests/test_gpu_guard.py -x -q > gpurun_out/try_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/try_tests.log
for m in 1 0 1 0; do
  printf "B2_WIDE_MERGE=%s: " $m; B2_WIDE_MERGE=$m timeout 120 python profiles/profile_wide.py 2>&1 | grep update_resident | tail -1
done | tee gpurun_out/try_ab.txt
for k in ${MASKS:-32 1}; do for m in 1 0; do
  printf "B2_WIDE_MERGE=%s skip %2d: " $m $k; B2_WIDE_MERGE=$m B2_WIDE_SKIP=$k timeout 120 python profiles/profile_wide.py 2>&1 | grep update_resident | tail -1
done; done | tee -a gpurun_out/try_ab.txt
