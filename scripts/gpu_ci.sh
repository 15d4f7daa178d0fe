#!/usr/bin/env bash
# Runs every GPU test file in its own process (a CUDA fault in one file must not poison the
# others) with a per-file timeout, and collects logs under gpurun_out/.
# Usage (on the GPU box, from the repo root):  bash scripts/gpu_ci.sh [file ...]
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/nvidia_smi.csv 2>&1
files=("$@")
if [ ${#files[@]} -eq 0 ]; then files=(tests/test_gpu_*.py tests/test_zz_*.py); fi
status=0
for f in "${files[@]}"; do
  [ -f "$f" ] || continue
  name=$(basename "$f" .py)
  echo "=== $f"
  timeout 600 python -m pytest "$f" -m gpu -q --tb=short -p no:cacheprovider > "gpurun_out/${name}.log" 2>&1
  rc=$?
  tail -n 25 "gpurun_out/${name}.log"
  echo "=== $f exit $rc"
  [ $rc -ne 0 ] && status=1
done
exit $status
