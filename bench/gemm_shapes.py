"""Times the GEMM shapes of one BERT-base training step (2,048 tokens) for every tile variant,
next to cuBLAS (torch.matmul) on the same operands.

    python bench/gemm_shapes.py [--tokens 2048] [--warm]

Each launch is timed alone with CUDA events (so every number carries the same ~5 us of launch and
event overhead, for cuBLAS too); L2 is flushed between launches unless --warm.  `loop` columns time
20 back-to-back launches and divide (launch overhead hidden, operands L2-warm).
Writes gpurun_out/gemm_shapes.json.
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dtds_b200.ops import kernels as K  # noqa: E402

PAIR = 1024


def timed(fn, flush, iters=15):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        else:
            torch.cuda._sleep(100_000)   # keep the GPU busy while the host queues the launch (hides host latency)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def looped(fn, n=20):
    fn()
    torch.cuda.synchronize()
    torch.cuda._sleep(200_000)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tokens", type=int, default=2048)
    ap.add_argument("--hidden", type=int, default=768)
    ap.add_argument("--ffn", type=int, default=3072)
    ap.add_argument("--warm", action="store_true")
    a = ap.parse_args()
    T, H, F = a.tokens, a.hidden, a.ffn
    dev = torch.device("cuda")
    flush = None if a.warm else torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    # (name, M, N, K, a_mn, b_mn): forward y = x W^T; dgrad dx = dy W (W MN-major); wgrad dW = dy^T x
    shapes = [("fwd qkv", T, 3 * H, H, 0, 0), ("fwd out", T, H, H, 0, 0), ("fwd fc1", T, F, H, 0, 0),
              ("fwd fc2", T, H, F, 0, 0),
              ("dgrad qkv", T, H, 3 * H, 0, 1), ("dgrad out", T, H, H, 0, 1), ("dgrad fc1", T, H, F, 0, 1),
              ("dgrad fc2", T, F, H, 0, 1),
              ("wgrad qkv", 3 * H, H, T, 1, 1), ("wgrad out", H, H, T, 1, 1), ("wgrad fc1", F, H, T, 1, 1),
              ("wgrad fc2", H, F, T, 1, 1), ("square 8192", 8192, 8192, 8192, 0, 0), ("square 4096", 4096, 4096, 4096, 0, 0)]
    variants = [("bn64", 64), ("bn128", 128), ("bn256", 256), ("pair128", PAIR + 128), ("pair256", PAIR + 256)]
    rows = []
    for name, M, N, Kd, a_mn, b_mn in shapes:
        A = torch.randn((Kd, M) if a_mn else (M, Kd), device=dev).to(torch.bfloat16)
        B = torch.randn((Kd, N) if b_mn else (N, Kd), device=dev).to(torch.bfloat16)
        out = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
        line = {"gemm": name, "M": M, "N": N, "K": Kd}
        fl = 2.0 * M * N * Kd
        for tag, bn in variants:
            f = lambda: K.gemm(A, B, a_mn_major=bool(a_mn), b_mn_major=bool(b_mn), out=out, block_n=bn)  # noqa: E731
            us, lp = timed(f, flush), looped(f)
            line[tag] = {"us": round(us, 1), "loop_us": round(lp, 2), "tflops_loop": round(fl / lp / 1e6)}
        Am = A.t() if a_mn else A
        Bm = B if b_mn else B.t()
        f = lambda: torch.matmul(Am, Bm, out=out)  # noqa: E731
        us, lp = timed(f, flush), looped(f)
        line["cublas"] = {"us": round(us, 1), "loop_us": round(lp, 2), "tflops_loop": round(fl / lp / 1e6)}
        rows.append(line)
        print(f"{name:12s} M={M:5d} N={N:5d} K={Kd:5d}  " + "  ".join(
            f"{t}: {line[t]['us']:6.1f}/{line[t]['loop_us']:7.2f}" for t in [v[0] for v in variants] + ["cublas"]), flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump({"tokens": T, "l2_flushed": not a.warm, "columns": "single-launch us / back-to-back loop us", "rows": rows},
              open(f"gpurun_out/gemm_shapes{'_warm' if a.warm else ''}.json", "w"), indent=1)


if __name__ == "__main__":
    main()
