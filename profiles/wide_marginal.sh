#!/bin/bash
# marginal in-pipeline cost of each kernel of the configs[4] optimizer step: time the update with one kernel left out
for m in 0 1 2 4 8 16 32 63; do
  printf "skip mask %2d: " $m; B2_WIDE_SKIP=$m python profiles/profile_wide.py 2>&1 | grep update_resident | tail -1
done
