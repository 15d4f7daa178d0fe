"""Times the GEMM shapes of one BERT-base training step (2,048 tokens) for every BLOCK_N.

    python bench/gemm_shapes.py [--tokens 2048] [--warm]

CUDA events, L2 flushed between launches unless --warm; prints TFLOP/s and the number of CTAs.
Writes gpurun_out/gemm_shapes.json.
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dtds_b200.ops import kernels as K  # noqa: E402


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
              ("wgrad fc2", H, F, T, 1, 1)]
    rows = []
    for name, M, N, Kd, a_mn, b_mn in shapes:
        A = torch.randn((Kd, M) if a_mn else (M, Kd), device=dev).to(torch.bfloat16)
        B = torch.randn((Kd, N) if b_mn else (N, Kd), device=dev).to(torch.bfloat16)
        out = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
        line = {"gemm": name, "M": M, "N": N, "K": Kd}
        for bn in (64, 128, 256):
            tiles = ((M + 127) // 128) * ((N + bn - 1) // bn)
            us = timed(lambda: K.gemm(A, B, a_mn_major=bool(a_mn), b_mn_major=bool(b_mn), out=out, block_n=bn), flush)
            line[f"bn{bn}"] = {"us": round(us, 1), "tflops": round(2.0 * M * N * Kd / us / 1e6), "tiles": tiles}
        rows.append(line)
        print(f"{name:10s} M={M:5d} N={N:5d} K={Kd:5d}  " + "  ".join(
            f"bn{bn}: {line[f'bn{bn}']['us']:6.1f} us {line[f'bn{bn}']['tflops']:4d} TF ({line[f'bn{bn}']['tiles']} tiles)"
            for bn in (64, 128, 256)))
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump({"tokens": T, "l2_flushed": not a.warm, "rows": rows},
              open(f"gpurun_out/gemm_shapes{'_warm' if a.warm else ''}.json", "w"), indent=1)


if __name__ == "__main__":
    main()
