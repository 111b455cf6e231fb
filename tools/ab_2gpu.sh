#!/bin/bash
# 2-GPU A/B of the gradient exchange: NVLink peer memory fused into reduce/Adam (B2_COMM_P2P=1, default) vs ncclAllReduce (=0)
timeout 600 python -m pytest tests/test_gpu_multi.py -q -m gpu -x 2>&1 | tail -6
for m in 1 0; do
  B2_COMM_P2P=$m timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2951$m \
    bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_2gpu_p2p$m.json 2> gpurun_out/bench_2gpu_p2p$m.err
  python -c "
import json; d=json.loads(open('gpurun_out/bench_2gpu_p2p$m.json').read().strip().splitlines()[-1])
print('P2P=$m rollout %.4g  update %.4g (%.2f ms)  wide %.4g (%.2f ms)' % (d['value'], d['update']['value'], d['update']['ms_per_update'], d['update_wide']['value'], d['update_wide']['ms_per_update']))"
done
