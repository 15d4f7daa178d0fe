#!/bin/bash
# Code written after the round-1 GPU budget was spent; run this first on the next GPU slot.
#   1 GPU : split-K GEMM numerics, then the step with and without it
#   2 GPUs: sparse-row DDP mode (add --gpus 2 to gpurun and pass "multi" as the first argument)
mkdir -p gpurun_out
if [ "$1" = "multi" ]; then
  DTDS_TEST_SPARSE_ROWS=1 python -m pytest tests/test_gpu_ops.py tests/test_gpu_multi.py -x -q -k "sparse or deterministic" 2>&1 | tail -30
  exit 0
fi
DTDS_TEST_SPLITK=1 DTDS_TEST_GEMM_WIDE=1 DTDS_TEST_PENDING=1 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_attention.py -q -k "splitk or wide or not_multiple" 2>&1 | tail -30
show() { python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', d['ms_per_step'], d['final_loss'])"; }
python bench.py --steps 20 --warmup 5 2>/dev/null | show baseline
DTDS_SPLITK=1 python bench.py --steps 20 --warmup 5 2>gpurun_out/splitk.err | show DTDS_SPLITK=1
DTDS_GEMM_WIDE=1 python bench.py --steps 20 --warmup 5 2>gpurun_out/wide.err | show DTDS_GEMM_WIDE=1
python bench/gemm_shapes.py > gpurun_out/gemm_shapes_plain.txt; DTDS_GEMM_WIDE=1 python bench/gemm_shapes.py > gpurun_out/gemm_shapes_wide.txt
paste gpurun_out/gemm_shapes_plain.txt gpurun_out/gemm_shapes_wide.txt | cut -c1-400
python bench/kernels_roofline.py | tee gpurun_out/kernels_roofline.txt   # copy the JSON to profiles/ afterwards
