#!/bin/bash
# A/B of programmatic dependent launch: full GPU tests with it on, then bench on and off.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
for v in 1 0; do
DTDS_PDL=$v timeout 300 python bench.py --steps 20 --warmup 5 2>gpurun_out/pdl_$v.err | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('PDL', $v, d['ms_per_step'], d['e2e']['ms_per_step'], d['final_loss'])"
done
DTDS_PDL=1 timeout 300 python bench.py --steps 20 --warmup 5 --no-graph 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('PDL 1 eager', d['ms_per_step'])"
DTDS_PDL=0 timeout 300 python bench.py --steps 20 --warmup 5 --no-graph 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('PDL 0 eager', d['ms_per_step'])"
tail -3 gpurun_out/pdl_1.err
