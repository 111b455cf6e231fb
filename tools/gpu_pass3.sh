#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_update_tc.py tests/test_gpu_guard.py -q -m gpu -s > gpurun_out/pytest_tc.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_tc.log
timeout 300 python tools/update_ab.py > gpurun_out/update_ab.log 2>&1
tail -40 gpurun_out/pytest_tc.log; cat gpurun_out/update_ab.log
