#!/bin/bash
# N-GPU A/B of the gradient exchange: fused flag-in-data exchange over NVLink peer memory (B2_COMM_P2P=1) vs ncclAllReduce (=0)
N=${1:-8}
for m in 1 0; do
  B2_COMM_P2P=$m timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$m \
    bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${N}gpu_p2p$m.json 2> gpurun_out/bench_${N}gpu_p2p$m.err
  tail -c 400 gpurun_out/bench_${N}gpu_p2p$m.err | grep -v "OMP_NUM\|\*\*\*"
  python -c "
import json; d=json.loads(open('gpurun_out/bench_${N}gpu_p2p$m.json').read().strip().splitlines()[-1])
print('N=$N P2P=$m rollout %.4g  e2e %.4g  update %.4g (%.2f ms)  wide %.4g (%.2f ms)' % (d['value'], d['e2e']['value'], d['update']['value'], d['update']['ms_per_update'], d['update_wide']['value'], d['update_wide']['ms_per_update']))"
done
