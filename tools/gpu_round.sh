#!/bin/bash
# One gpurun call (the round's final pass): full GPU test suite, smoke, both bench arms, launch lists, ncu --set full of the hot kernels
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_f64.json 2> gpurun_out/bench_f64.err; echo "bench rc=$?"
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-c4 --no-configs1 --no-wide-whole > gpurun_out/ncu_launch.log 2>&1
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/wide_launches.csv \
  python profiles/profile_wide.py 524288 32 1 1 > gpurun_out/wide_ncu.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"k_wide_fh|k_gemm|k_wide_l1r" -s 3 -c 3 -f -o gpurun_out/prof_wide_gemm \
  python profiles/profile_wide.py 524288 32 1 1 > gpurun_out/wide_ncu2.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_rollout -c 1 -f -o gpurun_out/prof_rollout_f64 \
  python profiles/profile_target.py f64 > gpurun_out/ncu_rollout_f64.log 2>&1
# %globaltimer timeline of one wide step: needs the diagnostic build (make TIMELINE=1) copied to ab_libs/timeline.so beforehand
if [ -f ab_libs/timeline.so ]; then
  L=reinforcement_learning_cpp_parallel_b200/libb200ppo.so; cp $L /tmp/lib_keep.so; cp ab_libs/timeline.so $L
  B2_WIDE_TL=1 timeout 120 python profiles/profile_wide.py 2>&1 | grep "^tl\|update_resident" > gpurun_out/wide_timeline.txt
  cp /tmp/lib_keep.so $L
fi
tail -3 gpurun_out/pytest_gpu.log; tail -1 gpurun_out/smoke.log
