#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_update_tc.py -q -m gpu -s -k large > gpurun_out/pytest_tc.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_tc.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_ppo_grad_tc -c 1 -f -o gpurun_out/prof_grad_tc \
  python profiles/profile_target.py f64 > gpurun_out/ncu_grad_tc.log 2>&1
tail -5 gpurun_out/pytest_tc.log; tail -3 gpurun_out/ncu_grad_tc.log
