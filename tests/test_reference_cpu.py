"""The fp32 oracle (ops/reference.py) against PyTorch itself.  The GPU numerics tests compare every
kernel with this module, so these tests close the chain kernel == oracle == torch on the CPU."""
import math

import pytest
import torch
import torch.nn.functional as F

from dtds_b200.ops import reference as ref
from dtds_b200.optim import AdamW, FusedAdam


def test_fused_adam_matches_torch_adamw_and_adam():
    torch.manual_seed(0)
    for adamw, wd in ((True, 0.01), (False, 0.0), (False, 0.1)):
        w0 = [torch.randn(37, 5), torch.randn(11), torch.randn(3, 4, 5)]
        a = [torch.nn.Parameter(t.clone()) for t in w0]
        b = [torch.nn.Parameter(t.clone()) for t in w0]
        mine = FusedAdam(a, lr=3e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=wd, adam_w_mode=adamw)
        cls = torch.optim.AdamW if adamw else torch.optim.Adam
        theirs = cls(b, lr=3e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=wd)
        for step in range(6):
            grads = [torch.randn_like(t) * (1 + step) for t in w0]
            mine.zero_grad()
            for p, g in zip(a, grads):
                p.grad = g.clone()
            for p, g in zip(b, grads):
                p.grad = g.clone()
            mine.step()
            theirs.step()
        for p, q in zip(a, b):
            torch.testing.assert_close(p.detach(), q.detach(), rtol=2e-6, atol=2e-7)


def test_layernorm_forward_backward_match_autograd():
    torch.manual_seed(1)
    x = torch.randn(6, 33, 96, requires_grad=True)
    res = torch.randn(6, 33, 96, requires_grad=True)
    g = torch.randn(96, requires_grad=True)
    b = torch.randn(96, requires_grad=True)
    y, z, mean, rstd = ref.layernorm_fwd(x.detach(), res.detach(), g.detach(), b.detach(), 1e-5)
    want = F.layer_norm(x + res, (96,), g, b, 1e-5)
    torch.testing.assert_close(y, want.detach(), rtol=1e-5, atol=1e-5)
    dy, extra = torch.randn_like(want), torch.randn_like(want)
    (want * dy).sum().backward()
    dz, dxd, dg, db = ref.layernorm_bwd(dy, z, mean, rstd, g.detach(), dz_extra=extra)
    torch.testing.assert_close(dz - extra, x.grad, rtol=1e-4, atol=1e-5)   # dz_extra is added on top
    torch.testing.assert_close(dg.reshape(-1), g.grad, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(db.reshape(-1), b.grad, rtol=1e-4, atol=1e-4)


def test_dropout_mask_is_a_pure_function_with_the_right_rate():
    m1 = ref.dropout_mask_scale((64, 1024), 0.1, 1234, torch.device("cpu"))
    m2 = ref.dropout_mask_scale((64, 1024), 0.1, 1234, torch.device("cpu"))
    m3 = ref.dropout_mask_scale((64, 1024), 0.1, 1235, torch.device("cpu"))
    assert torch.equal(m1, m2) and not torch.equal(m1, m3)
    kept = (m1 != 0).float().mean().item()
    assert abs(kept - 0.9) < 0.01
    assert torch.allclose(m1[m1 != 0], torch.tensor(1 / 0.9), rtol=1e-3)   # threshold is quantised to 1/65536
    assert abs(m1.mean().item() - 1.0) < 0.02                             # unbiased


@pytest.mark.parametrize("causal,alibi", [(False, False), (True, False), (True, True)])
def test_attention_matches_sdpa(causal, alibi):
    torch.manual_seed(2)
    B, h, S, d = 2, 3, 40, 16
    q, k, v = (torch.randn(B, h, S, d) for _ in range(3))
    slopes = torch.tensor([0.5, 0.25, 0.125]) if alibi else None
    kb = torch.zeros(B, S)
    kb[1, -7:] = -1e9
    out, lse = ref.attention(q, k, v, scale=1 / math.sqrt(d), causal=causal, key_bias=kb, alibi_slopes=slopes)
    bias = kb[:, None, None, :].expand(B, h, S, S).clone()
    if alibi:
        pos = torch.arange(S)
        bias = bias + slopes[None, :, None, None] * (pos[None, :] - pos[:, None])[None, None].float()
    if causal:
        bias = bias + torch.full((S, S), float("-inf")).triu(1)
    want = F.scaled_dot_product_attention(q, k, v, attn_mask=bias)
    torch.testing.assert_close(out, want, rtol=1e-4, atol=1e-5)
    scores = (q @ k.transpose(-1, -2)) / math.sqrt(d) + bias
    torch.testing.assert_close(lse, torch.logsumexp(scores, -1), rtol=1e-4, atol=1e-4)


def test_cross_entropy_shift_and_ignore_index():
    torch.manual_seed(3)
    logits = torch.randn(2 * 10, 50)
    labels = torch.randint(0, 50, (2 * 10,))
    labels[::4] = -100
    loss = ref.cross_entropy(logits, labels)
    loss = loss[0] if isinstance(loss, tuple) else loss
    torch.testing.assert_close(loss, F.cross_entropy(logits, labels, ignore_index=-100), rtol=1e-5, atol=1e-6)
    shifted = ref.cross_entropy(logits, labels, shift=True, seq=10)
    shifted = shifted[0] if isinstance(shifted, tuple) else shifted
    lg, lb = logits.view(2, 10, 50)[:, :-1].reshape(-1, 50), labels.view(2, 10)[:, 1:].reshape(-1)
    torch.testing.assert_close(shifted, F.cross_entropy(lg, lb, ignore_index=-100), rtol=1e-5, atol=1e-6)


def test_gemm_epilogues_and_activation_derivatives():
    torch.manual_seed(4)
    a, b, bias, res = torch.randn(9, 12), torch.randn(7, 12), torch.randn(7), torch.randn(9, 7)
    for act, fn in ((ref.ACT_RELU, F.relu), (ref.ACT_GELU_ERF, F.gelu),
                    (ref.ACT_GELU_TANH, lambda t: F.gelu(t, approximate="tanh"))):
        got = ref.gemm(a, b, bias=bias, act=act, residual=res)
        torch.testing.assert_close(got, fn(a @ b.t() + bias) + res, rtol=1e-5, atol=1e-5)
        x = torch.randn(200, requires_grad=True)
        fn(x).sum().backward()
        torch.testing.assert_close(ref.act_grad(x.detach(), act), x.grad, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(ref.gemm(a.t().contiguous(), b.t().contiguous(), a_mn_major=True, b_mn_major=True),
                               a @ b.t(), rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("causal", [False, True])
def test_attention_padded_to_block_multiple_matches_unpadded(causal):
    """Sequence lengths that are not a multiple of 128 run padded on the GPU kernels; the padding
    wrapper (pure autograd ops around the fused function) must not change values or gradients."""
    from dtds_b200.ops import functional as Fn
    torch.manual_seed(5)
    B, S, heads, d = 2, 100, 2, 64
    qkv = torch.randn(B * S, 3 * heads * d, requires_grad=True)
    kb = torch.zeros(B, S)
    kb[0, -9:] = -1e9
    want = Fn.AttentionFn.apply(qkv, B, S, heads, 0.125, causal, kb, None, 0.0, False)
    g = torch.randn_like(want)
    (want * g).sum().backward()
    gw, qkv.grad = qkv.grad.clone(), None
    got = Fn.attention_padded(qkv, B, S, heads, scale=0.125, causal=causal, key_bias=kb)
    (got * g).sum().backward()
    torch.testing.assert_close(got, want, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(qkv.grad, gw, rtol=1e-4, atol=1e-6)


def test_ddp_bucket_assignment_splits_the_tail():
    """Reverse-order buckets with a MiB cap; the last cap's worth of ordinary parameters is split so
    that only a small bucket is left at the very end of backward; oversize parameters sit alone."""
    import torch
    from dtds_b200.parallel.ddp import assign_buckets
    sizes = [30_000_000] + [2_000_000] * 40          # an oversize table first, then 40 ordinary weights
    params = [torch.nn.Parameter(torch.empty(0)) for _ in sizes]
    for p, n in zip(params, sizes):
        p.numel = (lambda n=n: n)                    # sizes only: no storage needed
    offs, o = [], 0
    for n in sizes:
        offs.append(o)
        o += n
    cap_mb, esz = 25, 2
    cap = cap_mb * 1024 * 1024 // esz
    buckets = assign_buckets(params, offs, cap_mb, esz)
    covered = sorted(i for idxs, _, _ in buckets for i in idxs)
    assert covered == list(range(len(sizes))), "every parameter in exactly one bucket"
    assert buckets[-1][0] == [0], "the oversize parameter is the last bucket, alone"
    dense = [sum(sizes[i] for i in idxs) for idxs, _, _ in buckets[:-1]]
    assert all(n <= cap for n in dense)
    assert dense[-1] <= cap // 2 + 2_000_000, f"tail bucket not split: {dense[-3:]}"
    for idxs, lo, hi in buckets:                     # contiguous ranges in the flat layout
        assert lo == min(offs[i] for i in idxs) and hi == max(offs[i] + sizes[i] for i in idxs)


def test_zero3_gradient_arena_layout():
    """Stage-3 gradients live in fixed space (tied / root parameters) plus a ring of block slots:
    buckets of one block never overlap, a block reuses exactly the slot of the block `depth` earlier in
    backward order, every parameter's main_grad is a view of the arena, and the arena is smaller than
    the model."""
    import torch
    from dtds_b200.models import build_model, get_config
    from dtds_b200.parallel import zero
    cfg = get_config("bloom-tiny").replace(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, num_hidden_layers=6)
    torch.manual_seed(0)
    engine, _, _, _ = zero.initialize(model=build_model(cfg), config={
        "train_micro_batch_size_per_gpu": 1, "optimizer": {"type": "Adam", "params": {"lr": 1e-3}},
        "zero_optimization": {"stage": 3, "reduce_bucket_size": 2e4, "grad_ring_blocks": 2}, "bf16": {"enabled": False}})
    m = engine.param3
    total = m.grads.numel()
    assert total < sum(p._numel for p in engine.flat.params), "arena as large as the model"
    spans = {}
    for b in engine.buckets:
        assert 0 <= b.goff and b.goff + (b.hi - b.lo) <= total
        spans[b.index] = (b.goff, b.goff + b.hi - b.lo)
        for p in b.params:
            o = b.goff + (p._flat_offset - b.lo)
            assert p.main_grad.data_ptr() == m.grads[o:o + 1].data_ptr() and p.main_grad.numel() == p._numel
    ring = [b for b in engine.buckets if b.prev_user is not None]
    assert ring, "no slot is ever reused: the test does not cover the ring"
    block_of = {id(q): i for i, blk in enumerate(m.blocks) for q in blk.params}
    for blk, waits in m._slot_waits.items():
        mine = [b for b in engine.buckets if b.prev_user is not None or b in ring
                if max(block_of.get(id(p), -1) for p in b.params) == blk]
        if waits and mine:    # a block starts at the slot base its predecessor `depth` blocks earlier used
            assert min(b.goff for b in waits) == min(b.goff for b in mine)
    # buckets that are live at the same time (same block group, or fixed) never overlap
    groups = {}
    for b in engine.buckets:
        key = ("fixed",) if b.prev_user is None and b.goff < min((x.goff for x in ring), default=total) else None
        groups.setdefault(key, []).append(b)
    fixed = groups.get(("fixed",), [])
    for i, a in enumerate(fixed):
        for c in fixed[i + 1:]:
            assert spans[a.index][1] <= spans[c.index][0] or spans[c.index][1] <= spans[a.index][0]
    # one training step runs and the loss is finite
    ids = torch.randint(0, cfg.vocab_size, (1, 64))
    loss = engine(ids, labels=ids).loss
    engine.backward(loss)
    engine.step()
    assert torch.isfinite(loss)
