#!/bin/bash
# Single-GPU check the driver repeats at round end: GPU tests, smoke(), default bench, estimator.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; tail -1 gpurun_out/bench_final.json | cut -c1-200
python bench.py --impl reference 2>/dev/null | tail -1 | cut -c1-200
python estimate_transformer_memory.py > gpurun_out/estimator_default.txt 2>&1; tail -12 gpurun_out/estimator_default.txt
