#!/bin/bash
# tools/sanitize.sh -- SURVEY.md s7.4 T11: compute-sanitizer over every hot kernel at small shapes.
#   memcheck   all four cases        racecheck  rollout / update / wide (shared-memory hazards)
#   synccheck  update / wide         initcheck  all four cases
# Summaries land in $OUT (default gpurun_out/) as sanitizer_<tool>.txt; copy them to profiles/ to commit.
# Usage (GPU box): bash tools/sanitize.sh [outdir] [tools...]
set -u
cd "$(dirname "$0")/.."
OUT=${1:-gpurun_out}
shift || true
TOOLS=${*:-memcheck racecheck synccheck initcheck}
mkdir -p "$OUT"
CS=${COMPUTE_SANITIZER:-/usr/local/cuda/bin/compute-sanitizer}
for tool in $TOOLS; do
  case $tool in
    memcheck|initcheck) cases="rollout update wide rollout_host" ;;
    racecheck)          cases="rollout update wide" ;;
    synccheck)          cases="update wide" ;;
    *) echo "unknown tool $tool"; exit 2 ;;
  esac
  f="$OUT/sanitizer_$tool.txt"
  : > "$f"
  for c in $cases; do
    echo "=== compute-sanitizer --tool $tool : sanitize_target.py $c ===" >> "$f"
    start=$(date +%s)
    # --print-limit keeps a pathological run readable; python's own allocations are host-side and not instrumented
    timeout 900 "$CS" --tool "$tool" --print-limit 20 --error-exitcode 77 \
      python tools/sanitize_target.py "$c" > "$OUT/.san_${tool}_${c}.log" 2>&1
    rc=$?
    grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize_target|Error:|Hazard|=========     at |Invalid|Uninitialized|Barrier error" "$OUT/.san_${tool}_${c}.log" | head -60 >> "$f"
    echo "exit code $rc, $(( $(date +%s) - start )) s" >> "$f"
  done
done
tail -n +1 "$OUT"/sanitizer_*.txt
