#!/bin/bash
# One gpurun call: parity tests, both bench arms, ncu launch list, ncu --set full of the hot kernels.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_f64.json 2> gpurun_out/bench_f64.err; echo "bench rc=$?"
timeout 600 python bench.py --precision f32 --no-cpu-baseline > gpurun_out/bench_f32.json 2> gpurun_out/bench_f32.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
for p in f64 f32; do
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_rollout -c 1 -f -o gpurun_out/prof_rollout_$p \
  python profiles/profile_target.py $p > gpurun_out/ncu_rollout_$p.log 2>&1
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_ppo_grad -c 1 -f -o gpurun_out/prof_grad \
  python profiles/profile_target.py f64 > gpurun_out/ncu_grad.log 2>&1
# configs[4] (2x256 update sweep on tensor cores): launch list + full captures of the GEMM and the side kernels
timeout 300 python profiles/profile_wide.py > gpurun_out/wide_timing.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/wide_launches.csv \
  python profiles/profile_wide.py 524288 32 1 1 > gpurun_out/wide_ncu.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_gemm -s 3 -c 3 -f -o gpurun_out/prof_wide_gemm \
  python profiles/profile_wide.py 524288 32 1 1 > gpurun_out/wide_ncu2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_wide_l1|k_wide_head|k_wide_dw1|k_wide_reduce" -s 8 -c 4 -f \
  -o gpurun_out/prof_wide_side python profiles/profile_wide.py 524288 32 1 1 > gpurun_out/wide_ncu3.log 2>&1
ls -la gpurun_out
