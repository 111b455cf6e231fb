#!/bin/bash
# 2-GPU pass: multi-GPU parity tests + the bench line with its multi_gpu_check, C4 strong-scaling leg at N = 2
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -q -m gpu -x > gpurun_out/pytest_multi.log 2>&1; echo "pytest_multi rc=$?" | tee -a gpurun_out/pytest_multi.log
timeout 600 python bench.py --no-wide-whole > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; echo "bench1 rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; echo "bench2 rc=$?"
tail -3 gpurun_out/pytest_multi.log; tail -c 1500 gpurun_out/bench_2gpu.err
