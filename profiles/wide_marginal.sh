#!/bin/bash
# marginal in-pipeline cost of each kernel of the configs[4] optimizer step: time the update with one kernel left out
# (1 l1, 2|4 fwd GEMM + head (6 = the fused k_wide_fh), 8 dX GEMM, 16 dW2 GEMM + dw1, 32 reduce, 64 dW2 GEMM without its dw1 warps' work)
for m in ${MASKS:-0 1 6 8 16 32 64 7 63}; do
  printf "skip mask %2d: " $m; B2_WIDE_SKIP=$m python profiles/profile_wide.py 2>&1 | grep update_resident | tail -1
done
