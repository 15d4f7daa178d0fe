#!/bin/bash
# `ncu --set full` captures of one launch per single-GPU kernel family (1 GPU; never multi-rank).
# The plain program runs first and must exit 0; every capture is a separate ncu process on the same
# command line (bench/kernels_roofline.py --iters 2), filtered to one kernel name.
mkdir -p gpurun_out
CMD="python bench/kernels_roofline.py --iters 2"
$CMD > gpurun_out/ncu_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu_plain.log; exit 1; }
for k in ln_fwd_warp_kernel ln_bwd_dx_kernel ln_bwd_dgb_kernel ce_fwd_bwd_kernel adam_flat_kernel embedding_fwd_kernel colsum_kernel attn_fwd_kernel attn_bwd_dkdv_kernel attn_bwd_dq_kernel; do
  timeout 150 ncu --set full --clock-control none --import-source on -k regex:$k -s 2 -c 1 -f -o gpurun_out/ncu_$k $CMD > gpurun_out/ncu_$k.log 2>&1
  echo "$k rc=$? $(ls -la gpurun_out/ncu_$k.ncu-rep 2>/dev/null | awk '{print $5}')"
done
