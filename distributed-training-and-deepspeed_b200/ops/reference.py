"""Plain PyTorch reference implementations of every op in ``csrc/``.

Two jobs: (1) the oracle the GPU numerics tests compare each CUDA kernel against (computed in
fp32), (2) the implementation used for CPU tensors, which is what the gloo/CPU plumbing tests and
the estimators run.  Nothing here is used for CUDA tensors.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

ACT_NONE, ACT_RELU, ACT_GELU_ERF, ACT_GELU_TANH = 0, 1, 2, 3
ACT_BY_NAME = {None: ACT_NONE, "none": ACT_NONE, "relu": ACT_RELU, "gelu": ACT_GELU_ERF,
               "gelu_erf": ACT_GELU_ERF, "gelu_tanh": ACT_GELU_TANH, "gelu_new": ACT_GELU_TANH}


def act_fwd(x: torch.Tensor, act: int) -> torch.Tensor:
    if act == ACT_RELU:
        return F.relu(x)
    if act == ACT_GELU_ERF:
        return F.gelu(x)
    if act == ACT_GELU_TANH:
        return F.gelu(x, approximate="tanh")
    return x


def act_grad(x: torch.Tensor, act: int) -> torch.Tensor:
    """d act(x) / dx evaluated elementwise."""
    xf = x.float()
    if act == ACT_RELU:
        return (xf > 0).float()
    if act == ACT_GELU_ERF:
        cdf = 0.5 * (1.0 + torch.erf(xf / math.sqrt(2.0)))
        pdf = torch.exp(-0.5 * xf * xf) / math.sqrt(2.0 * math.pi)
        return cdf + xf * pdf
    if act == ACT_GELU_TANH:
        k = math.sqrt(2.0 / math.pi)
        u = k * (xf + 0.044715 * xf ** 3)
        t = torch.tanh(u)
        du = k * (1.0 + 3 * 0.044715 * xf * xf)
        return 0.5 * (1 + t) + 0.5 * xf * (1 - t * t) * du
    return torch.ones_like(xf)


def gemm(a, b, *, a_mn_major=False, b_mn_major=False, bias=None, act=ACT_NONE, residual=None,
         dact_in=None, alpha=1.0, out_dtype=None, accumulate_into=None, want_pre=False):
    """D = act(alpha * A·Bᵀ + bias) (+ residual).  A is [M,K] (or stored [K,M] if a_mn_major),
    B is [N,K] (or stored [K,N] if b_mn_major)."""
    af = a.float().t() if a_mn_major else a.float()
    bf = b.float().t() if b_mn_major else b.float()
    d = alpha * (af @ bf.t())
    if bias is not None:
        d = d + bias.float()
    pre = d
    if dact_in is not None:
        d = d * act_grad(dact_in, act)
    else:
        d = act_fwd(d, act)
    if residual is not None:
        d = d + residual.float()
    if accumulate_into is not None:
        d = d + accumulate_into.float()
    out_dtype = out_dtype or a.dtype
    if want_pre:
        return d.to(out_dtype), pre.to(a.dtype)
    return d.to(out_dtype)


def _mix32(h: torch.Tensor) -> torch.Tensor:
    """murmur3 fmix32 on int64 tensors holding uint32 values."""
    m = 0xFFFFFFFF
    h = h ^ (h >> 16)
    h = (h * 0x85EBCA6B) & m
    h = h ^ (h >> 13)
    h = (h * 0xC2B2AE35) & m
    h = h ^ (h >> 16)
    return h


def dropout_mask_scale(shape, p: float, seed: int, device) -> torch.Tensor:
    """Keep-scale tensor (0 or 1/(1-p)) for the counter-based dropout of csrc/common.cuh:
    16 random bits per element, two elements per 32-bit hash of (seed, element_index / 2)."""
    if p <= 0.0:
        return torch.ones(shape, device=device)
    n = 1
    for s in shape:
        n *= s
    m = 0xFFFFFFFF
    idx = torch.arange(n, dtype=torch.int64, device=device)
    pair = idx >> 1
    seed &= (1 << 64) - 1
    seed_lo, seed_hi = seed & m, (seed >> 32) & m
    pair_lo, pair_hi = pair & m, (pair >> 32) & m
    key = seed_lo ^ _mix32((seed_hi + ((pair_hi * 0x85EBCA77) & m)) & m)
    bits = _mix32(key ^ ((pair_lo * 0x9E3779B1) & m))
    u = torch.where((idx & 1) == 1, bits >> 16, bits & 0xFFFF)
    thr = int(p * 65536.0)
    keep = (u >= thr).float() / (1.0 - p)
    return keep.view(shape)


def layernorm_fwd(x, residual, gamma, beta, eps, p_drop=0.0, seed=0):
    """z = dropout(x) + residual ; y = LN(z).  Returns (y, z, mean, rstd)."""
    xf = x.float()
    if p_drop > 0.0:
        xf = xf * dropout_mask_scale(x.shape, p_drop, seed, x.device)
    if residual is not None:
        xf = xf + residual.float()
    z = xf.to(x.dtype)
    zf = z.float()
    mean = zf.mean(-1)
    var = zf.var(-1, unbiased=False)
    rstd = torch.rsqrt(var + eps)
    y = (zf - mean[..., None]) * rstd[..., None] * gamma.float() + beta.float()
    return y.to(x.dtype), z, mean, rstd


def layernorm_bwd(dy, z, mean, rstd, gamma, dz_extra=None, p_drop=0.0, seed=0):
    """Returns (dz, dx_drop, dgamma, dbeta) in fp32."""
    dyf, zf = dy.float(), z.float()
    xhat = (zf - mean[..., None]) * rstd[..., None]
    g = dyf * gamma.float()
    s1 = g.mean(-1, keepdim=True)
    s2 = (g * xhat).mean(-1, keepdim=True)
    dz = rstd[..., None] * (g - s1 - xhat * s2)
    if dz_extra is not None:
        dz = dz + dz_extra.float()
    dz = dz.to(dy.dtype).float() if dy.dtype != torch.float32 else dz
    dx_drop = dz * dropout_mask_scale(z.shape, p_drop, seed, z.device) if p_drop > 0 else dz
    cols = z.shape[-1]
    dgamma = (dyf * xhat).reshape(-1, cols).sum(0)
    dbeta = dyf.reshape(-1, cols).sum(0)
    return dz, dx_drop, dgamma, dbeta


def attention(q, k, v, *, scale, causal=False, key_bias=None, alibi_slopes=None, p_drop=0.0,
              seed=0):
    """q,k,v: [B, H, S, D].  Returns (out [B,H,S,D], lse [B,H,S]) in fp32.
    key_bias: [B, S] additive (padding mask); alibi_slopes: [H] -> bias slope*(j - i)."""
    B, H, S, D = q.shape
    s = torch.einsum("bhid,bhjd->bhij", q.float(), k.float()) * scale
    if alibi_slopes is not None:
        pos = torch.arange(S, device=q.device)
        rel = (pos[None, :] - pos[:, None]).float()
        s = s + alibi_slopes.float().view(1, H, 1, 1) * rel.view(1, 1, S, S)
    if key_bias is not None:
        s = s + key_bias.float().view(B, 1, 1, S)
    if causal:
        m = torch.ones(S, S, dtype=torch.bool, device=q.device).tril()
        s = s.masked_fill(~m, float("-inf"))
    lse = torch.logsumexp(s, dim=-1)
    p = torch.exp(s - lse[..., None])
    if p_drop > 0.0:
        p = p * dropout_mask_scale((B, H, S, S), p_drop, seed, q.device)
    out = torch.einsum("bhij,bhjd->bhid", p, v.float())
    return out, lse


def cross_entropy(logits, labels, *, ignore_index=-100, shift=False, seq=None):
    """Mean CE over non-ignored rows.  logits [R, V]; labels [R].  shift: causal-LM convention."""
    lf = logits.float()
    lab = labels.clone()
    if shift:
        lab = lab.view(-1, seq)
        lab = torch.cat([lab[:, 1:], torch.full_like(lab[:, :1], ignore_index)], dim=1).reshape(-1)
    return F.cross_entropy(lf, lab, ignore_index=ignore_index), lab


def ce_valid_count(labels, ignore_index=-100, shift=False, seq=None):
    """Number of rows that count towards the mean in `cross_entropy` (0-dim tensor)."""
    lab = labels
    if shift:
        lab = lab.view(-1, seq)[:, 1:]
    return (lab != ignore_index).sum()


def adam_step(p, g, m, v, step, *, lr, beta1, beta2, eps, weight_decay, adamw_mode=True,
              bias_correction=True, grad_scale=1.0):
    """One fp32 Adam/AdamW update in place; `step` is the 1-based step number."""
    g = g.float() * grad_scale
    if not adamw_mode and weight_decay != 0.0:
        g = g + weight_decay * p
    m.mul_(beta1).add_(g, alpha=1 - beta1)
    v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
    c1 = 1 - beta1 ** step if bias_correction else 1.0
    c2 = 1 - beta2 ** step if bias_correction else 1.0
    denom = v.sqrt() / math.sqrt(c2) + eps
    if adamw_mode and weight_decay != 0.0:
        p.mul_(1 - lr * weight_decay)
    p.addcdiv_(m, denom, value=-lr / c1)
    return p
