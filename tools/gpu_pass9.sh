#!/bin/bash
# round-2 re-entry sanity pass: GPU parity tests, smoke, the default bench line
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_f64.json 2> gpurun_out/bench_f64.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/bench_f64.json
tail -5 gpurun_out/pytest_gpu.log
