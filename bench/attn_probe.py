"""Fused attention forward/backward at the bench shape (BERT-base: B=4, S=512, 12 heads, d=64).

    python bench/attn_probe.py [--p-drop 0.1] [--iters 20]

CUDA events on the launching stream, L2 flushed between launches, median.  Prints TFLOP/s against
the 4*B*heads*S^2*d (forward) and 10*B*heads*S^2*d (backward: 5 GEMMs) FLOP counts.
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dtds_b200.ops import kernels as K  # noqa: E402


def timed(fn, iters, flush):
    ts = []
    for _ in range(3):
        fn()
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--seq", type=int, default=512)
    ap.add_argument("--heads", type=int, default=12)
    ap.add_argument("--p-drop", type=float, default=0.1)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--causal", action="store_true")
    a = ap.parse_args()
    B, S, h, d = a.batch, a.seq, a.heads, 64
    dev = torch.device("cuda")
    qkv = (torch.randn(B * S, 3 * h * d, device=dev) * 0.5).to(torch.bfloat16)
    dout = torch.randn(B * S, h * d, device=dev).to(torch.bfloat16)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    kw = dict(scale=d ** -0.5, causal=a.causal, p_drop=a.p_drop, seed=1234)
    out, lse = K.attention_fwd(qkv, B, S, h, **kw)
    rows = []
    for p in sorted({0.0, a.p_drop}):
        kw["p_drop"] = p
        t_f = timed(lambda: K.attention_fwd(qkv, B, S, h, **kw), a.iters, flush)
        t_b = timed(lambda: K.attention_bwd(qkv, out, dout, lse, B, S, h, **kw), a.iters, flush)
        fl = B * h * S * S * d * (0.5 if a.causal else 1.0)
        rows.append({"p_drop": p, "fwd_us": round(t_f, 1), "fwd_tflops": round(4 * fl / t_f / 1e6, 1),
                     "bwd_us": round(t_b, 1), "bwd_tflops": round(10 * fl / t_b / 1e6, 1)})
        print(rows[-1])
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump({"B": B, "S": S, "heads": h, "d": d, "causal": a.causal, "rows": rows},
              open("gpurun_out/attn_probe.json", "w"), indent=1)


if __name__ == "__main__":
    main()
