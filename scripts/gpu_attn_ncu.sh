#!/bin/bash
# ncu --set full of the dropout-specialised attention kernels at the bench shape (run after the
# same program exited 0 without ncu).
mkdir -p gpurun_out
python bench/attn_probe.py --iters 1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:attn_ -s 19 -c 6 -o gpurun_out/attn_ncu2 -f python bench/attn_probe.py --iters 1 > gpurun_out/attn_ncu2.log 2>&1
tail -2 gpurun_out/attn_ncu2.log
