"""Fused tcgen05 attention forward/backward vs an fp32 PyTorch reference."""
import json
import math
import os

import pytest
import torch

from dtds_b200.ops import kernels as K
from dtds_b200.ops import reference as ref

pytestmark = pytest.mark.gpu
DEV = "cuda"
OUT_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")


def _stats(name, got, want):
    got, want = got.float(), want.float()
    err = (got - want).abs()
    return {"what": name, "max_err": err.max().item(), "mean_err": err.mean().item(),
            "ref_max": want.abs().max().item(), "ref_mean": want.abs().mean().item(),
            "nan": bool(torch.isnan(got).any().item())}


def _ok(s, rel):
    return (not s["nan"]) and s["max_err"] <= rel * s["ref_max"] + 1e-3 and \
        s["mean_err"] <= 0.25 * rel * s["ref_max"] + 1e-4


CASES = [
    dict(B=1, S=128, heads=1),
    dict(B=2, S=512, heads=12),
    dict(B=2, S=256, heads=4, causal=True),
    dict(B=2, S=256, heads=4, key_bias=True),
    dict(B=1, S=384, heads=16, causal=True, alibi=True),
    dict(B=2, S=256, heads=2, p_drop=0.1),
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: "_".join(f"{k}{v}" for k, v in c.items()))
def test_attention_fwd_bwd(case):
    torch.manual_seed(0)
    B, S, heads = case["B"], case["S"], case["heads"]
    d, H = 64, case["heads"] * 64
    causal = case.get("causal", False)
    p_drop = case.get("p_drop", 0.0)
    seed = 4242
    qkv = (torch.randn(B * S, 3 * H, device=DEV) * 0.8).to(torch.bfloat16)
    key_bias = None
    if case.get("key_bias"):
        key_bias = torch.zeros(B, S, device=DEV)
        key_bias[:, S - 37:] = -1e9  # padded tail
    alibi = None
    if case.get("alibi"):
        alibi = torch.tensor([2.0 ** (-8.0 * (i + 1) / heads) for i in range(heads)], device=DEV)
    scale = 1.0 / math.sqrt(d)
    kw = dict(scale=scale, causal=causal, key_bias=key_bias, alibi=alibi, p_drop=p_drop, seed=seed)

    out, lse = K.attention_fwd(qkv, B, S, heads, **kw)
    torch.cuda.synchronize()
    qkv_f = qkv.float().requires_grad_(True)
    q, k, v = K._split_qkv(qkv_f, B, S, heads)
    o_ref, lse_ref = ref.attention(q, k, v, scale=scale, causal=causal, key_bias=key_bias,
                                   alibi_slopes=alibi, p_drop=p_drop, seed=seed)
    o_ref2 = o_ref.transpose(1, 2).reshape(B * S, H)
    report = [_stats("out", out, o_ref2), _stats("lse", lse, lse_ref)]

    dout = (torch.randn(B * S, H, device=DEV) * 0.5).to(torch.bfloat16)
    dqkv = K.attention_bwd(qkv, out, dout, lse, B, S, heads, **kw)
    torch.cuda.synchronize()
    (g,) = torch.autograd.grad(o_ref2, qkv_f, dout.float())
    report += [_stats("dq", dqkv[:, :H], g[:, :H]), _stats("dk", dqkv[:, H:2 * H], g[:, H:2 * H]),
               _stats("dv", dqkv[:, 2 * H:], g[:, 2 * H:])]
    os.makedirs(OUT_DIR, exist_ok=True)
    tag = "_".join(f"{k}{v}" for k, v in case.items())
    with open(os.path.join(OUT_DIR, f"attn_{tag}.json"), "w") as f:
        json.dump(report, f, indent=1)
    bad = [s for s in report if not _ok(s, 0.03)]
    assert not bad, json.dumps(report, indent=1)


def test_attention_throughput_report():
    B, S, heads = 4, 512, 12
    H = heads * 64
    qkv = (torch.randn(B * S, 3 * H, device=DEV) * 0.5).to(torch.bfloat16)
    dout = (torch.randn(B * S, H, device=DEV) * 0.5).to(torch.bfloat16)
    kw = dict(scale=0.125)
    out, lse = K.attention_fwd(qkv, B, S, heads, **kw)
    K.attention_bwd(qkv, out, dout, lse, B, S, heads, **kw)
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    it = 20
    ev[0].record()
    for _ in range(it):
        out, lse = K.attention_fwd(qkv, B, S, heads, **kw)
    ev[1].record()
    for _ in range(it):
        K.attention_bwd(qkv, out, dout, lse, B, S, heads, **kw)
    ev[2].record()
    torch.cuda.synchronize()
    fwd_ms = ev[0].elapsed_time(ev[1]) / it
    bwd_ms = ev[1].elapsed_time(ev[2]) / it
    fl = 4.0 * B * heads * S * S * 64
    rep = {"B": B, "S": S, "heads": heads, "fwd_ms": fwd_ms, "fwd_tflops": fl / fwd_ms / 1e9,
           "bwd_ms": bwd_ms, "bwd_tflops": 2.5 * fl / bwd_ms / 1e9}
    with open(os.path.join(OUT_DIR, "attn_throughput.json"), "w") as f:
        json.dump(rep, f)
    print(rep)


@pytest.mark.parametrize("S", [100, 200, 333])
def test_attention_sequence_length_not_multiple_of_128(S):
    """Fn.attention pads to the kernels' 128-row block (masked keys, dropped query rows)."""
    import math
    from dtds_b200.ops import functional as Fn
    from dtds_b200.ops import reference as ref
    torch.manual_seed(0)
    B, heads, d = 2, 4, 64
    qkv = (torch.randn(B * S, 3 * heads * d, device="cuda") * 0.5).to(torch.bfloat16).requires_grad_(True)
    out = Fn.attention(qkv, B, S, heads, causal=True)
    g = torch.randn_like(out)
    (out.float() * g.float()).sum().backward()
    q32 = qkv.detach().float().cpu().requires_grad_(True)
    q, k, v = (t.reshape(B, S, heads, d).transpose(1, 2) for t in q32.split(heads * d, dim=-1))
    want, _ = ref.attention(q, k, v, scale=1 / math.sqrt(d), causal=True)
    want = want.transpose(1, 2).reshape(B * S, heads * d)
    (want * g.float().cpu()).sum().backward()
    assert (out.float().cpu() - want).abs().max().item() < 3e-2
    assert (qkv.grad.float().cpu() - q32.grad).abs().max().item() < 6e-2 * max(1.0, q32.grad.abs().max().item())
