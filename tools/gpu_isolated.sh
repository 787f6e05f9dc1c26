#!/usr/bin/env bash
# Run each GPU test of a file in its own process with a timeout, so one trapped/hung kernel does
# not hide the results of the others.  Usage: tools/gpu_isolated.sh tests/test_kernels_gpu.py [outfile]
set -u
FILE=${1:-tests/test_kernels_gpu.py}
OUT=${2:-gpurun_out/isolated.log}
mkdir -p "$(dirname "$OUT")"
: > "$OUT"
nvidia-smi --query-gpu=name,driver_version,memory.total --format=csv >> "$OUT" 2>&1
IDS=$(python -m pytest "$FILE" -m gpu --collect-only -q 2>/dev/null | grep '::')
for id in $IDS; do
  echo "=== $id" >> "$OUT"
  timeout 180 python -m pytest "$id" -m gpu -x -q --no-header -p no:cacheprovider 2>&1 | tail -n 25 >> "$OUT"
  echo "--- exit ${PIPESTATUS[0]}" >> "$OUT"
done
grep -E '^(===|--- exit)' "$OUT" | paste - - | awk '{print $NF, $2}' | sort | uniq -c | sort -rn | head -80
