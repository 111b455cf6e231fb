#!/bin/bash
# N-GPU end-to-end leg: rollout kernel writing straight into pinned host memory (default) vs the chunked copy-engine pipeline
N=${1:-8}
for m in 1 0; do
  B2_ROLLOUT_HOST_DIRECT=$m timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$m \
    bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline --no-wide > gpurun_out/bench_${N}gpu_direct$m.json 2> gpurun_out/bench_${N}gpu_direct$m.err
  python -c "
import json; d=json.loads(open('gpurun_out/bench_${N}gpu_direct$m.json').read().strip().splitlines()[-1])
print('N=$N host-direct=$m rollout %.4g  e2e %.4g  update %.4g' % (d['value'], d['e2e']['value'], d['update']['value']))"
done
