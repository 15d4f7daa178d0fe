#!/bin/bash
# Per-kernel device time of ~1.1 training steps (eager launches, warm caches).
mkdir -p gpurun_out
python bench.py --no-graph --steps 2 --warmup 6 > gpurun_out/step_plain.log 2>&1 || { tail -5 gpurun_out/step_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -s 2600 -c 420 --csv \
    --log-file gpurun_out/step_launches.csv python bench.py --no-graph --steps 2 --warmup 6 > gpurun_out/step_ncu.log 2>&1
tail -2 gpurun_out/step_ncu.log | cut -c1-200
wc -l gpurun_out/step_launches.csv
