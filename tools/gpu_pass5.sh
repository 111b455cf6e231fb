#!/bin/bash
# 2-GPU: exchange tests incl. the forced scatter shape, bench with either shape
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -q -m gpu -x > gpurun_out/pytest_multi.log 2>&1; echo "pytest_multi rc=$?" | tee -a gpurun_out/pytest_multi.log
for sc in 0 1; do
B2_COMM_SCATTER=$sc timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2951$sc \
  bench.py --gpus 2 --steps 5 --warmup 3 --no-c4 --no-e2e > gpurun_out/bench_2gpu_sc$sc.json 2> gpurun_out/bench_2gpu_sc$sc.err; echo "bench sc=$sc rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/bench_2gpu_sc$sc.json').read().strip().splitlines()[-1])
print('scatter=$sc update %.4g (%.2f ms)  wide %.4g (%.2f ms) check', d['update']['value'], d['update']['ms_per_update'], d['update_wide']['value'], d['update_wide']['ms_per_update'], d['multi_gpu_check']['ok'], d['multi_gpu_check']['p2p_vs_nccl_relerr'])"
done
tail -4 gpurun_out/pytest_multi.log
