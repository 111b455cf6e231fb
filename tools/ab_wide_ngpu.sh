#!/bin/bash
# 8-GPU A/B of the wide step: fused forward + head (default) vs the two kernels of round 1, NVLink byte counters
set -u
mkdir -p gpurun_out
N=${1:-8}
for f in 1 0; do
  B2_WIDE_FUSED=$f timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2955$f \
    bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline --no-c4 --no-configs1 --no-e2e --no-update --no-multi-check \
    > gpurun_out/bench_${N}gpu_fused$f.json 2> gpurun_out/bench_${N}gpu_fused$f.err
  python -c "
import json; d=json.loads(open('gpurun_out/bench_${N}gpu_fused$f.json').read().strip().splitlines()[-1]); w=d['update_wide']
print('N=$N fused=$f wide %.4g sample-epochs/s  %.2f ms per update (max over ranks)  frac %.3f' % (w['value'], w['ms_per_update'], w['roofline']['frac'])); print(w['roofline'].get('nvlink'))"
done
