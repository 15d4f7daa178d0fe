"""Achieved fraction of roofline for every single-GPU kernel family at the bench shapes.

    python bench/kernels_roofline.py            # writes gpurun_out/kernels_roofline.json

Denominators come from MEASURED_PEAKS.json (copy bandwidth, cuBLAS bf16 burst throughput); each
row's roofline time is max(bytes / HBM bandwidth, FLOPs / bf16 peak).  CUDA events on the launching
stream, L2 flushed between launches, median of `--iters`.
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dtds_b200.ops import kernels as K  # noqa: E402
from dtds_b200.ops.reference import ACT_GELU_ERF  # noqa: E402

try:
    _P = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    HBM_GBS, PEAK_TF = float(_P["hbm_gbs"]), float(_P["bf16_tflops"])
except Exception:  # noqa: BLE001
    HBM_GBS, PEAK_TF = 6500.0, 1590.0


def timed(fn, flush, iters):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
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
    ap.add_argument("--vocab", type=int, default=28996)
    ap.add_argument("--iters", type=int, default=15)
    a = ap.parse_args()
    T, H, V = a.tokens, a.hidden, a.vocab
    dev = torch.device("cuda")
    bf = torch.bfloat16
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    rows = []

    def add(name, us, nbytes, flops=0.0):
        t_mem = nbytes / (HBM_GBS * 1e9) * 1e6
        t_cmp = flops / (PEAK_TF * 1e12) * 1e6
        roof = max(t_mem, t_cmp)
        rows.append({"kernel": name, "time_us": round(us, 2), "roofline_us": round(roof, 2),
                     "fraction": round(roof / us, 3), "bound": "compute" if t_cmp > t_mem else "hbm",
                     "bytes": int(nbytes), "flops": float(flops)})
        print(f"{name:44s} {us:8.1f} us  roofline {roof:7.2f} us  {roof / us:5.2f}  ({rows[-1]['bound']})")

    def rnd(*s, scale=1.0):
        return (torch.randn(*s, device=dev) * scale).to(bf)

    # ---- GEMMs of one BERT-base layer (forward shapes) + the vocabulary projection ---------------
    for name, M, N, Kd in [("gemm qkv", T, 3 * H, H), ("gemm out-proj", T, H, H), ("gemm ffn up", T, 4 * H, H),
                           ("gemm ffn down", T, H, 4 * H), ("gemm vocab", T, V // 8 * 8, H), ("gemm 8192^3", 8192, 8192, 8192)]:
        x, w = rnd(M, Kd), rnd(N, Kd, scale=0.05)
        out = torch.empty(M, N, dtype=bf, device=dev)
        us = timed(lambda: K.gemm(x, w, out=out), flush, a.iters)
        add(f"{name} {M}x{N}x{Kd}", us, 2.0 * (M * Kd + N * Kd + M * N), 2.0 * M * N * Kd)
    x, w, bias, res = rnd(T, H), rnd(4 * H, H, scale=0.05), rnd(4 * H), rnd(T, 4 * H)
    pre = torch.empty(T, 4 * H, dtype=bf, device=dev)
    out = torch.empty(T, 4 * H, dtype=bf, device=dev)
    us = timed(lambda: K.gemm(x, w, bias=bias, act=ACT_GELU_ERF, residual=res, pre_out=pre, out=out), flush, a.iters)
    add("gemm ffn up + bias + GELU + residual + pre", us, 2.0 * (T * H + 4 * H * H + 3 * T * 4 * H), 2.0 * T * 4 * H * H)

    # ---- attention (B x 512 tokens, 12 heads, d = 64) --------------------------------------------
    S, heads = 512, H // 64
    B = max(T // S, 1)
    qkv, dout = rnd(B * S, 3 * H, scale=0.5), rnd(B * S, H)
    for p in (0.0, 0.1):
        kw = dict(scale=0.125, p_drop=p, seed=7)
        o, lse = K.attention_fwd(qkv, B, S, heads, **kw)
        fl = 4.0 * B * heads * S * S * 64
        us = timed(lambda: K.attention_fwd(qkv, B, S, heads, **kw), flush, a.iters)
        add(f"attention fwd p_drop={p}", us, 2.0 * B * S * 4 * H, fl)
        us = timed(lambda: K.attention_bwd(qkv, o, dout, lse, B, S, heads, **kw), flush, a.iters)
        add(f"attention bwd p_drop={p}", us, 2.0 * B * S * 8 * H, 2.5 * fl)

    # ---- LayerNorm, reductions, loss, embedding, optimizer ---------------------------------------
    xx, rr, g, bb = rnd(T, H), rnd(T, H), rnd(H), rnd(H)
    y, z, mean, rstd = K.layernorm_fwd(xx, rr, g, bb, 1e-5, 0.1, 3)
    us = timed(lambda: K.layernorm_fwd(xx, rr, g, bb, 1e-5, 0.1, 3), flush, a.iters)
    add("layernorm fwd (+residual +dropout)", us, 2.0 * T * H * 4)
    dy = rnd(T, H)
    us = timed(lambda: K.layernorm_bwd(dy, z, mean, rstd, g, p_drop=0.1, seed=3, want_dx_drop=True, parts="dx"),
               flush, a.iters)
    add("layernorm bwd dz (+dropout)", us, 2.0 * T * H * 4)
    dg, dbt = torch.zeros(H, dtype=bf, device=dev), torch.zeros(H, dtype=bf, device=dev)
    us = timed(lambda: K.layernorm_bwd(dy, z, mean, rstd, g, dgamma_out=dg, dbeta_out=dbt, parts="params"),
               flush, a.iters)
    add("layernorm bwd dgamma, dbeta", us, 2.0 * T * H * 2)
    wide = rnd(T, 4 * H)
    us = timed(lambda: K.colsum(wide), flush, a.iters)
    add("column sum (bias gradient) 2048x3072", us, 2.0 * T * 4 * H)
    Vp = (V + 7) // 8 * 8
    logits = rnd(T, Vp)
    labels = torch.randint(0, V, (T,), device=dev)
    labels[::7] = -100
    us = timed(lambda: K.cross_entropy(logits[:, :V], labels, vocab=V), flush, a.iters)
    add("cross-entropy fwd+bwd in place", us, 2.0 * T * Vp * 2)
    word, pos, typ = rnd(V, H), rnd(512, H), rnd(2, H)
    ids = torch.randint(0, V, (B, S), device=dev)
    us = timed(lambda: K.embedding_fwd(ids, word, pos, typ, None, 0), flush, a.iters)
    add("embedding gather", us, 2.0 * B * S * H * 2)
    n = 108_000_000 // 8 * 8
    p32, g16 = torch.zeros(n, device=dev), torch.zeros(n, dtype=bf, device=dev)
    m, v = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    plp = torch.empty(n, dtype=bf, device=dev)
    step = torch.zeros((), dtype=torch.int32, device=dev)
    us = timed(lambda: K.adam_flat(p32, g16, m, v, step, lr=1e-4, p_lp=plp), flush, max(a.iters // 3, 3))
    add("fused AdamW, 108 M parameters", us, n * 28.0)

    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "kernels_roofline.json"), "w") as f:
        json.dump({"hbm_GBps": HBM_GBS, "bf16_peak_TFLOPs": PEAK_TF, "tokens": T, "hidden": H, "rows": rows}, f, indent=1)


if __name__ == "__main__":
    main()
