#!/bin/bash
# Code written after the round-1 GPU budget was spent; run this first on the next GPU slot.
#   1 GPU : split-K GEMM numerics, then the step with and without it
#   2 GPUs: sparse-row DDP mode (add --gpus 2 to gpurun and pass "multi" as the first argument)
mkdir -p gpurun_out
if [ "$1" = "multi" ]; then
  DTDS_TEST_SPARSE_ROWS=1 python -m pytest tests/test_gpu_ops.py tests/test_gpu_multi.py -x -q -k "sparse or deterministic" 2>&1 | tail -30
  exit 0
fi
DTDS_TEST_SPLITK=1 python -m pytest tests/test_gpu_gemm.py -x -q -k splitk 2>&1 | tail -30
for v in 0 1; do
  DTDS_SPLITK=$v python bench.py --steps 20 --warmup 5 2>gpurun_out/splitk_$v.err | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('DTDS_SPLITK', $v, d['ms_per_step'], d['final_loss'])"
done
