#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_rollout.py tests/test_gpu_variants.py -q -m gpu -x > gpurun_out/pytest_roll.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_roll.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py --no-wide --no-c4 --no-configs1 --no-cpu-baseline --steps 10 > gpurun_out/bench_roll.json 2> gpurun_out/bench_roll.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/bench_roll.json').read().strip().splitlines()[-1])
print('rollout %.4g (%.3f ms) frac %.3f  e2e %.4g  update %.4g' % (d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'], d['update']['value']))"
tail -4 gpurun_out/pytest_roll.log
