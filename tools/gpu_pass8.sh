#!/bin/bash
# e2e A/B (direct host stores vs copy-engine pipeline) + the round's ncu captures
set -u
mkdir -p gpurun_out
for m in 1 0; do for c in 8 16; do
B2_ROLLOUT_HOST_DIRECT=$m python bench.py --no-cpu-baseline --no-wide --no-c4 --no-configs1 --no-update --steps 8 --warmup 3 --chunks $c 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('direct=$m chunks=$c value %.4g e2e %.4g'%(d['value'], d['e2e']['value']))"; done; done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-c4 --no-configs1 --no-wide-whole > gpurun_out/ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_rollout -c 1 -f -o gpurun_out/prof_rollout_f64 \
  python profiles/profile_target.py f64 > gpurun_out/ncu_rollout_f64.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_ppo_grad_tc -c 1 -f -o gpurun_out/prof_grad_tc \
  python profiles/profile_target.py f64 > gpurun_out/ncu_grad_tc.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/wide_launches.csv \
  python profiles/profile_wide.py 524288 32 1 1 > gpurun_out/wide_ncu.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_gemm -s 3 -c 3 -f -o gpurun_out/prof_wide_gemm \
  python profiles/profile_wide.py 524288 32 1 1 > gpurun_out/wide_ncu2.log 2>&1
ls -la gpurun_out | head -30
