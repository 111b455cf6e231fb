#!/bin/bash
# marginal in-pipeline cost of each kernel of the configs[4] optimizer step: time the update with one kernel left out
# (1 l1, 2 fwd GEMM, 4 head, 8 dX GEMM, 16 dW2 GEMM + dw1, 32 reduce, 64 dW2 GEMM without its dw1 warps' work)
for m in 0 1 2 4 8 16 32 64 3 6 12 7 63; do
  printf "skip mask %2d: " $m; B2_WIDE_SKIP=$m python profiles/profile_wide.py 2>&1 | grep update_resident | tail -1
done
