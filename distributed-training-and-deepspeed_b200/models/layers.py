"""Fused building blocks shared by the BERT and causal-LM models.

Every module is a thin parameter holder; the math is in ``ops.functional`` (hand-written sm_100a
kernels on CUDA, fp32 reference ops on CPU).  Activations are kept token-major ``[B, S, H]`` /
``[B*S, H]`` end to end: the QKV projection output is consumed by the attention kernel in place
and the context comes back token-major, so there are no head-major transposes or ``contiguous``
copies anywhere.
"""
from __future__ import annotations

import math

import torch
from torch import nn

from ..ops import ACT_BY_NAME, ACT_NONE
from ..ops import functional as Fn


def _init_normal_(t, std=0.02):
    with torch.no_grad():
        t.normal_(0.0, std)
    return t


class Linear(nn.Module):
    """y = act(x·Wᵀ + b) (+ residual) as one GEMM with a fused epilogue."""

    def __init__(self, in_features, out_features, bias=True, act=None, init_std=0.02):
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.act = ACT_BY_NAME[act]
        self.weight = nn.Parameter(_init_normal_(torch.empty(out_features, in_features), init_std))
        self.bias = nn.Parameter(torch.zeros(out_features)) if bias else None

    def forward(self, x, residual=None, fork=False):
        return Fn.linear(x, self.weight, self.bias, self.act, residual, fork)

    def extra_repr(self):
        return f"in={self.in_features}, out={self.out_features}, bias={self.bias is not None}, act={self.act}"


class LayerNorm(nn.Module):
    """y = LN(dropout(x) + residual) — the residual add and the dropout in front of it are part
    of the kernel (BERT: ``LayerNorm(dropout(dense(x)) + input)``)."""

    def __init__(self, hidden, eps=1e-5, p_drop=0.0):
        super().__init__()
        self.eps, self.p_drop = eps, p_drop
        self.weight = nn.Parameter(torch.ones(hidden))
        self.bias = nn.Parameter(torch.zeros(hidden))

    def forward(self, x, residual=None, fork=False):
        return Fn.layer_norm(x, self.weight, self.bias, self.eps, residual, self.p_drop, self.training, fork)


class MLP(nn.Module):
    def __init__(self, hidden, ffn, act="gelu", bias=True):
        super().__init__()
        self.act = ACT_BY_NAME[act]
        self.fc1 = Linear(hidden, ffn, bias=bias)
        self.fc2 = Linear(ffn, hidden, bias=bias)

    def forward(self, x, residual=None, fork=False):
        return Fn.mlp(x, self.fc1.weight, self.fc1.bias, self.fc2.weight, self.fc2.bias, self.act, residual, fork)


def alibi_slopes(num_heads: int) -> torch.Tensor:
    """ALiBi head slopes as in BLOOM (closest power of two + interleaved remainder)."""
    closest = 2 ** math.floor(math.log2(num_heads))
    base = 2.0 ** (-(2.0 ** -(math.log2(closest) - 3)))
    slopes = [base ** (i + 1) for i in range(closest)]
    if closest != num_heads:
        extra_base = 2.0 ** (-(2.0 ** -(math.log2(2 * closest) - 3)))
        n_extra = min(closest, num_heads - closest)
        slopes += [extra_base ** (2 * i + 1) for i in range(n_extra)]
    return torch.tensor(slopes, dtype=torch.float32)


class SelfAttention(nn.Module):
    """Fused QKV projection -> fused attention kernel -> output projection."""

    def __init__(self, hidden, heads, p_drop=0.0, causal=False, alibi=False, bias=True):
        super().__init__()
        assert hidden % heads == 0
        self.hidden, self.heads, self.p_drop, self.causal = hidden, heads, p_drop, causal
        self.qkv = Linear(hidden, 3 * hidden, bias=bias)
        self.out = Linear(hidden, hidden, bias=bias)
        # ZeRO-3 may leave this weight un-gathered and let the QKV GEMM pull it from the owners'
        # shards itself (all-gather fused with the first consuming GEMM of the block): it is only
        # ever read through Fn.linear, which knows how to do that.
        self.qkv.weight._ag_ok = True
        # fp32 slopes, kept out of the buffer registry so `.to(bfloat16)` cannot down-cast them
        # (the attention kernel reads them as float*)
        self._alibi_slopes = alibi_slopes(heads) if alibi else None
        self._alibi_by_device = {}

    def alibi_on(self, device):
        if self._alibi_slopes is None:
            return None
        t = self._alibi_by_device.get(device)
        if t is None:
            t = self._alibi_slopes.to(device=device, dtype=torch.float32)
            self._alibi_by_device[device] = t
        return t

    def forward(self, x, key_bias=None, residual=None, fork=False):
        """``fork=True`` returns ``(out, x_skip)`` (see ops.functional.linear)."""
        B, S, H = x.shape
        x_skip = None
        if fork:
            qkv, x_skip = self.qkv(x, fork=True)
            qkv = qkv.view(B * S, 3 * H)
        else:
            qkv = self.qkv(x).view(B * S, 3 * H)
        ctx = Fn.attention(qkv, B, S, self.heads, causal=self.causal, key_bias=key_bias,
                           alibi=self.alibi_on(x.device), p_drop=self.p_drop, training=self.training)
        out = self.out(ctx.view(B, S, H), residual)
        return (out, x_skip) if fork else out


class Dropout(nn.Module):
    """Standalone dropout (+ residual) with a recomputable mask."""

    def __init__(self, p):
        super().__init__()
        self.p = p

    def forward(self, x, residual=None):
        if not self.training or self.p <= 0.0:
            return x if residual is None else x + residual
        return _DropoutFn.apply(x, residual, self.p)


class _DropoutFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, residual, p):
        from ..ops import kernels as K
        ctx.seed, ctx.p, ctx.has_res = Fn.next_seed(), p, residual is not None
        return K.dropout(x, p, ctx.seed, residual)

    @staticmethod
    def backward(ctx, dy):
        from ..ops import kernels as K
        return K.dropout(dy, ctx.p, ctx.seed), (dy if ctx.has_res else None), None


def key_padding_bias(attention_mask: torch.Tensor | None) -> torch.Tensor | None:
    """[B, S] {0,1} mask -> additive fp32 bias (0 keep, -30000 drop) for the attention kernel."""
    if attention_mask is None:
        return None
    return (1.0 - attention_mask.float()) * -30000.0


__all__ = ["Linear", "LayerNorm", "MLP", "SelfAttention", "Dropout", "alibi_slopes",
           "key_padding_bias", "ACT_NONE"]
