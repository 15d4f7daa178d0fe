#!/bin/bash
mkdir -p gpurun_out
python bench/attn_probe.py || exit 1
ncu --set full --clock-control none --import-source on -k regex:attn_ -s 12 -c 4 -o gpurun_out/attn_ncu -f python bench/attn_probe.py --iters 2 > gpurun_out/attn_ncu.log 2>&1
tail -3 gpurun_out/attn_ncu.log
ls -la gpurun_out/attn_ncu.ncu-rep
