"""Runs one GEMM shape a few times (for `ncu --set full` captures and quick CUDA-event timings).

    python bench/gemm_probe.py M N K [a_mn] [b_mn] [block_n] [iters]
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dtds_b200.ops import kernels as K  # noqa: E402

M, N, Kd = (int(x) for x in sys.argv[1:4])
a_mn = bool(int(sys.argv[4])) if len(sys.argv) > 4 else False
b_mn = bool(int(sys.argv[5])) if len(sys.argv) > 5 else False
bn = int(sys.argv[6]) if len(sys.argv) > 6 else 0
iters = int(sys.argv[7]) if len(sys.argv) > 7 else 20
a = torch.randn((Kd, M) if a_mn else (M, Kd), device="cuda").to(torch.bfloat16)
b = torch.randn((Kd, N) if b_mn else (N, Kd), device="cuda").to(torch.bfloat16)
out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(3):
    K.gemm(a, b, a_mn_major=a_mn, b_mn_major=b_mn, out=out, block_n=bn)
torch.cuda.synchronize()
times = []
WARM = bool(os.environ.get("DTDS_WARM"))   # keep operands in L2 (a spin kernel hides host latency instead)
for _ in range(iters):
    if WARM:
        torch.cuda._sleep(100_000)
    else:
        flush.zero_()  # evict L2 between timed launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    K.gemm(a, b, a_mn_major=a_mn, b_mn_major=b_mn, out=out, block_n=bn)
    e1.record()
    torch.cuda.synchronize()
    times.append(e0.elapsed_time(e1) * 1e3)
times.sort()
med = times[len(times) // 2]
print(f"M={M} N={N} K={Kd} a_mn={int(a_mn)} b_mn={int(b_mn)} bn={bn}: median {med:.1f} us, min {times[0]:.1f} us, "
      f"{2.0 * M * N * Kd / med / 1e6:.1f} TFLOP/s ({'operands warm in L2' if WARM else 'L2 flushed between launches'})")

if os.environ.get("DTDS_GEMM_TIMELINE"):
    from dtds_b200.ops import _lib
    dbg = torch.zeros(148 * 16, dtype=torch.int64, device="cuda")
    _lib.lib().dtds_gemm_set_debug_buffer(dbg.data_ptr())
    if WARM:
        torch.cuda._sleep(100_000)
    else:
        flush.zero_()
    K.gemm(a, b, a_mn_major=a_mn, b_mn_major=b_mn, out=out, block_n=bn)
    torch.cuda.synchronize()
    _lib.lib().dtds_gemm_set_debug_buffer(None)
    t = dbg.view(148, 16).cpu()
    labels = ["setup", "loads_issued", "first_landed", "mma_issued", "acc_ready", "tile_stored", "end"]
    for cta in (0, 1, 50, 95):
        if t[cta, 0] == 0:
            continue
        rel = (t[cta, 1:8] - t[cta, 0]).tolist()
        epi = [(int(x) - int(t[cta, 0])) if x else 0 for x in t[cta, 8:14].tolist()]
        print(f"   epilogue warp2: chunk0 ld={epi[0]} math={epi[1]} stored={epi[2]} | chunk1 ld={epi[3]} math={epi[4]} stored={epi[5]}")
        print(f"CTA {cta}: " + "  ".join(f"{l}={v}" for l, v in zip(labels, rel)) + " (SM clocks from CTA start)")
