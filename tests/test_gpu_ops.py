"""Memory-bound sm_100a kernels (LayerNorm, activations, dropout, softmax, column sums, MSE,
embedding, cross-entropy, Adam) vs fp32 PyTorch references."""
import pytest
import torch
import torch.nn.functional as F

from dtds_b200.ops import kernels as K
from dtds_b200.ops import reference as ref

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _bf(*shape, scale=1.0):
    return (torch.randn(*shape, device=DEV) * scale).to(torch.bfloat16)


def _close(got, want, atol, rtol, what):
    got, want = got.float(), want.float()
    err = (got - want).abs()
    lim = atol + rtol * want.abs()
    assert not torch.isnan(got).any(), f"{what}: NaN"
    assert (err <= lim).all(), f"{what}: max err {err.max().item():.4g} (ref max {want.abs().max().item():.4g})"


@pytest.mark.parametrize("cols", [64, 256, 768, 1024, 9216])
@pytest.mark.parametrize("p_drop", [0.0, 0.1])
def test_layernorm_fwd_bwd(cols, p_drop):
    torch.manual_seed(0)
    rows = 300
    x, res = _bf(rows, cols), _bf(rows, cols)
    gamma, beta = _bf(cols, scale=0.5) + 1, _bf(cols, scale=0.1)
    seed = 1234
    y, z, mean, rstd = K.layernorm_fwd(x, res, gamma, beta, 1e-5, p_drop, seed)
    yr, zr, mr, rr = ref.layernorm_fwd(x, res, gamma, beta, 1e-5, p_drop, seed)
    _close(z, zr, 1e-2, 1e-2, "z")
    _close(mean, mr, 2e-3, 1e-3, "mean")
    _close(rstd, rr, 2e-3, 2e-3, "rstd")
    _close(y, yr, 3e-2, 2e-2, "y")
    dy = _bf(rows, cols)
    extra = _bf(rows, cols)
    dz, dxd, dg, db = K.layernorm_bwd(dy, z, mean, rstd, gamma, dz_extra=extra, p_drop=p_drop,
                                      seed=seed, want_dx_drop=True)
    dzr, dxdr, dgr, dbr = ref.layernorm_bwd(dy, z, mean, rstd, gamma, dz_extra=extra, p_drop=p_drop,
                                            seed=seed)
    _close(dz, dzr, 3e-2, 2e-2, "dz")
    _close(dxd, dxdr, 3e-2, 2e-2, "dx_drop")
    _close(dg, dgr, 0.25, 2e-2, "dgamma")
    _close(db, dbr, 0.25, 2e-2, "dbeta")
    # fp32 accumulate-into outputs
    dgo = torch.ones(cols, device=DEV)
    dbo = torch.ones(cols, device=DEV)
    K.layernorm_bwd(dy, z, mean, rstd, gamma, dgamma_out=dgo, dbeta_out=dbo, accumulate=True)
    _, _, dgr2, dbr2 = ref.layernorm_bwd(dy, z, mean, rstd, gamma)
    _close(dgo, dgr2 + 1, 5e-3, 1e-3, "dgamma fp32 accumulate")
    _close(dbo, dbr2 + 1, 5e-3, 1e-3, "dbeta fp32 accumulate")


def test_layernorm_matches_torch_layer_norm():
    x = _bf(64, 768)
    g, b = _bf(768) + 1, _bf(768)
    y, _, _, _ = K.layernorm_fwd(x, None, g, b, 1e-12)
    want = F.layer_norm(x.float(), (768,), g.float(), b.float(), 1e-12)
    _close(y, want, 3e-2, 2e-2, "layer_norm")


@pytest.mark.parametrize("act", [ref.ACT_RELU, ref.ACT_GELU_ERF, ref.ACT_GELU_TANH])
def test_activations(act):
    x = _bf(1000, 333, scale=2.0)  # odd size exercises the scalar tail
    y = K.act_fwd(x, act)
    _close(y, ref.act_fwd(x.float(), act), 1e-2, 1e-2, "act fwd")
    dy = _bf(1000, 333)
    dx = K.act_bwd(dy, x, act)
    xf = x.float().requires_grad_(True)
    ref.act_fwd(xf, act).backward(dy.float())
    _close(dx, xf.grad, 2e-2, 2e-2, "act bwd")


def test_dropout_matches_reference_hash_and_rate():
    x = torch.ones(512, 1024, device=DEV, dtype=torch.bfloat16)
    y = K.dropout(x, 0.1, 99)
    want = ref.dropout_mask_scale(x.shape, 0.1, 99, DEV)
    _close(y, want, 1e-2, 1e-2, "dropout mask")
    rate = (y == 0).float().mean().item()
    assert abs(rate - 0.1) < 0.01, rate
    r = _bf(512, 1024)
    y2 = K.dropout(x, 0.1, 99, residual=r)
    _close(y2, want + r.float(), 2e-2, 1e-2, "dropout+residual")


def test_softmax_rows_fwd_bwd():
    x = _bf(96, 512, scale=3.0)
    y = K.softmax_rows(x)
    _close(y, torch.softmax(x.float(), -1), 2e-3, 2e-2, "softmax")
    dy = _bf(96, 512)
    dx = K.softmax_rows_bwd(y, dy)
    yf = y.float()
    want = yf * (dy.float() - (yf * dy.float()).sum(-1, keepdim=True))
    _close(dx, want, 2e-3, 2e-2, "softmax bwd")


@pytest.mark.parametrize("rows,cols", [(2048, 768), (100, 70), (4096, 28996)])
def test_colsum(rows, cols):
    ld = (cols + 7) // 8 * 8
    x = _bf(rows, ld)[:, :cols]
    out = K.colsum(x)
    _close(out, x.float().sum(0), 0.5, 2e-2, "colsum")
    o32 = torch.ones(cols, device=DEV)
    K.colsum(x, out=o32, accumulate=True)
    _close(o32, x.float().sum(0) + 1, 2e-2, 1e-3, "colsum fp32 accumulate")
    # the shared scratch must be left zeroed: a second call gives the same answer
    _close(K.colsum(x), x.float().sum(0), 0.5, 2e-2, "colsum second call")


def test_mse():
    x, t = _bf(4, 512, 96), _bf(4, 512, 96)
    loss, dx = K.mse(x, t)
    xf = x.float().requires_grad_(True)
    want = F.mse_loss(xf, t.float())
    want.backward()
    _close(loss, want, 1e-3, 1e-3, "mse loss")
    _close(dx, xf.grad, 1e-7, 2e-2, "mse grad")


def test_embedding_fwd_bwd():
    torch.manual_seed(0)
    B, S, H, V = 4, 128, 768, 1000
    ids = torch.randint(0, V, (B, S), device=DEV)
    ids[0, :8] = 7  # repeated ids -> colliding scatter-adds
    word, pos, typ = _bf(V, H), _bf(512, H), _bf(2, H)
    out = K.embedding_fwd(ids, word, pos, typ)
    want = word[ids].float() + pos[:S].float()[None] + typ[0].float()
    _close(out, want, 3e-2, 1e-2, "embedding fwd")
    dout = _bf(B, S, H)
    for dt, atol in ((torch.float32, 1e-3), (torch.bfloat16, 0.15)):
        dword = torch.zeros(V, H, device=DEV, dtype=dt)
        dpos = torch.zeros(512, H, device=DEV, dtype=dt)
        K.embedding_bwd(ids, dout, dword, dpos)
        want_w = torch.zeros(V, H, device=DEV).index_add_(0, ids.reshape(-1), dout.reshape(-1, H).float())
        _close(dword, want_w, atol, 2e-2, f"dword {dt}")
        _close(dpos[:S], dout.float().sum(0), atol, 2e-2, f"dpos {dt}")


@pytest.mark.parametrize("V,ld", [(28996, 29056), (1000, 1000), (250880, 250880), (517, 520)])
@pytest.mark.parametrize("shift", [False, True])
def test_cross_entropy(V, ld, shift):
    torch.manual_seed(0)
    B, S = 2, 64
    R = B * S
    logits = torch.zeros(R, ld, device=DEV, dtype=torch.bfloat16)
    logits[:, :V] = _bf(R, V, scale=2.0)
    labels = torch.randint(0, V, (R,), device=DEV)
    labels[::7] = -100
    lf = logits[:, :V].float().requires_grad_(True)
    want, _ = ref.cross_entropy(lf, labels, shift=shift, seq=S)
    want.backward()
    loss, d = K.cross_entropy(logits.clone(), labels, vocab=V, shift=shift, seq=S,
                              inplace_grad=True)
    _close(loss, want.detach(), 2e-3, 2e-3, "ce loss")
    _close(d[:, :V], lf.grad, 2e-5, 3e-2, "ce grad")
    loss2, none = K.cross_entropy(logits.clone(), labels, vocab=V, shift=shift, seq=S, want_grad=False)
    assert none is None
    _close(loss2, want.detach(), 2e-3, 2e-3, "ce loss (no grad)")


@pytest.mark.parametrize("grad_dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("adamw,wd", [(True, 0.0), (True, 0.01), (False, 0.01)])
def test_adam_flat_matches_torch(grad_dtype, adamw, wd):
    torch.manual_seed(0)
    n = 100_003
    p0 = torch.randn(n, device=DEV)
    p = p0.clone()
    m, v = torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
    p_lp = torch.empty(n, device=DEV, dtype=torch.bfloat16)
    step = torch.zeros((), dtype=torch.int32, device=DEV)
    tp = torch.nn.Parameter(p0.clone())
    cls = torch.optim.AdamW if adamw else torch.optim.Adam
    opt = cls([tp], lr=1e-2, betas=(0.9, 0.999), eps=1e-8, weight_decay=wd)
    for _ in range(5):
        g = torch.randn(n, device=DEV).to(grad_dtype)
        K.adam_flat(p, g, m, v, step, lr=1e-2, eps=1e-8, weight_decay=wd, adamw_mode=adamw, p_lp=p_lp)
        tp.grad = g.float()
        opt.step()
    assert int(step.item()) == 5
    _close(p, tp.detach(), 2e-6, 2e-5, "adam params")
    _close(p_lp, tp.detach(), 1e-2, 1e-2, "bf16 copy")


def test_cast_roundtrip():
    x = torch.randn(12345, device=DEV)
    lp = torch.empty(12345, device=DEV, dtype=torch.bfloat16)
    K.cast(x, lp)
    assert torch.equal(lp, x.to(torch.bfloat16))
    back = torch.empty_like(x)
    K.cast(lp, back)
    assert torch.equal(back, lp.float())


