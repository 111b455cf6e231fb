#!/bin/bash
# 8-GPU: gradient exchange shapes A/B (wide + default update legs only), then the full bench line once
set -u
N=${1:-8}
mkdir -p gpurun_out
for sc in 1 0; do
B2_COMM_SCATTER=$sc timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$sc \
  bench.py --gpus $N --steps 5 --warmup 3 --no-c4 --no-e2e --no-cpu-baseline > gpurun_out/bench_${N}gpu_sc$sc.json 2> gpurun_out/bench_${N}gpu_sc$sc.err; echo "bench sc=$sc rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/bench_${N}gpu_sc$sc.json').read().strip().splitlines()[-1])
print('N=$N scatter=$sc update %.4g (%.2f ms)  wide %.4g (%.2f ms) frac %.3f check' % (d['update']['value'], d['update']['ms_per_update'], d['update_wide']['value'], d['update_wide']['ms_per_update'], d['update_wide']['roofline']['frac']), json.dumps(d['multi_gpu_check']))"
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29530 \
  bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_${N}gpu_full.json 2> gpurun_out/bench_${N}gpu_full.err; echo "bench full rc=$?"
tail -c 300 gpurun_out/bench_${N}gpu_full.err
