"""``TransformerBlock``: one OPT-style pre-LN decoder block in which *every* operation is its own
``nn.Module`` so forward hooks can observe each activation (the design goal stated at
model/transformer.py:19-26 of the reference; the memory estimators depend on it).

Sub-module attribute names — and therefore ``state_dict`` keys — match the reference
(model/transformer.py:36-53): ``pre_attention_layer_norm, k_proj, v_proj, q_proj, out_proj,
compute_attention_weights, attention_weights_softmax, attention_weights_dropout,
compute_attentions, attention_output_dropout, final_layer_norm, ffn_1, ffn_2, activation,
final_dropout``.  As in the reference there is no attention mask and no KV cache.

Two execution modes:
  * ``fused=False`` (default, what the estimators use): op-per-module.  bf16 CUDA tensors run the
    standalone sm_100a kernels (GEMM, LayerNorm, row-softmax, dropout, ReLU); fp32 tensors run
    the matching torch ops (the reference's numerics, used for like-for-like memory readings).
  * ``fused=True``: same parameters, but attention goes through the fused tcgen05 kernel (no
    [B·heads, S, S] tensor) when head_dim == 64 and S % 128 == 0.
"""
from __future__ import annotations

import torch
from torch import nn

from ..ops import ACT_NONE, ACT_RELU
from ..ops import functional as Fn
from ..ops import kernels as K


def _native(x: torch.Tensor) -> bool:
    return x.is_cuda and x.dtype == torch.bfloat16


class MatMul(nn.Module):
    """Module wrapper for a batched matmul so hooks see its output (model/transformer.py:6-15).
    Batched library GEMM: this unfused path exists for observability, not speed."""

    def forward(self, x, y):
        return torch.bmm(x, y)


class Linear(nn.Linear):
    def forward(self, x):
        if _native(x):
            return Fn.linear(x, self.weight, self.bias, ACT_NONE)
        return super().forward(x)


class LayerNorm(nn.LayerNorm):
    def forward(self, x):
        if _native(x) and x.shape[-1] % 8 == 0:
            return Fn.layer_norm(x, self.weight, self.bias, self.eps)
        return super().forward(x)


class _SoftmaxFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        y = K.softmax_rows(x)
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        return K.softmax_rows_bwd(y, dy)


class Softmax(nn.Module):
    def __init__(self, dim=-1):
        super().__init__()
        self.dim = dim

    def forward(self, x):
        if _native(x) and self.dim in (-1, x.dim() - 1):
            return _SoftmaxFn.apply(x)
        return torch.softmax(x, self.dim)


class _ActFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, act):
        ctx.save_for_backward(x)
        ctx.act = act
        return K.act_fwd(x, act)

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        return K.act_bwd(dy.contiguous(), x, ctx.act), None


class ReLU(nn.Module):
    def forward(self, x):
        if _native(x):
            return _ActFn.apply(x, ACT_RELU)
        return torch.relu(x)


class Dropout(nn.Module):
    """Class name must stay ``Dropout``: the activation counter charges one byte per element for
    modules of that name (estimate_nn_memory.py:31-33 in the reference)."""

    def __init__(self, p=0.0):
        super().__init__()
        self.p = p

    def forward(self, x):
        if not self.training or self.p <= 0.0:
            return x
        if _native(x):
            from .layers import _DropoutFn
            return _DropoutFn.apply(x, None, self.p)
        return torch.nn.functional.dropout(x, self.p, True)


class TransformerBlock(nn.Module):
    """forward(hidden_states [B, S, H]) -> [B, S, H]."""

    def __init__(self, config, fused: bool = False):
        super().__init__()
        self.config = config
        self.fused = fused
        self.hidden_size = config.hidden_size
        self.num_attention_heads = config.num_attention_heads
        self.head_dim = config.hidden_size // config.num_attention_heads
        self.ffn_dim = config.ffn_dim
        self.scaling = self.head_dim ** -0.5
        bias = getattr(config, "enable_bias", True)
        p = getattr(config, "dropout", 0.0)
        H = self.hidden_size

        self.pre_attention_layer_norm = LayerNorm(H)
        self.k_proj = Linear(H, H, bias=bias)
        self.v_proj = Linear(H, H, bias=bias)
        self.q_proj = Linear(H, H, bias=bias)
        self.out_proj = Linear(H, H, bias=bias)

        self.compute_attention_weights = MatMul()
        self.attention_weights_softmax = Softmax(dim=-1)
        self.attention_weights_dropout = Dropout(p)
        self.compute_attentions = MatMul()

        self.attention_output_dropout = Dropout(p)
        self.final_layer_norm = LayerNorm(H)

        self.ffn_1 = Linear(H, self.ffn_dim, bias=bias)
        self.ffn_2 = Linear(self.ffn_dim, H, bias=bias)
        self.activation = ReLU()
        self.final_dropout = Dropout(p)

    def _attention_unfused(self, x, B, S):
        heads, d = self.num_attention_heads, self.head_dim

        def split(t):
            return t.view(B, S, heads, d).transpose(1, 2).reshape(B * heads, S, d)

        q = split(self.q_proj(x) * self.scaling)
        k = split(self.k_proj(x))
        v = split(self.v_proj(x))
        w = self.compute_attention_weights(q, k.transpose(1, 2))
        w = self.attention_weights_softmax(w)
        w = self.attention_weights_dropout(w)
        o = self.compute_attentions(w, v)
        return o.view(B, heads, S, d).transpose(1, 2).reshape(B, S, self.hidden_size)

    def _attention_fused(self, x, B, S):
        qkv = torch.cat([self.q_proj(x), self.k_proj(x), self.v_proj(x)], dim=-1)
        p = self.attention_weights_dropout.p
        ctx = Fn.attention(qkv.view(B * S, -1), B, S, self.num_attention_heads, scale=self.scaling,
                           p_drop=p, training=self.training)
        return ctx.view(B, S, self.hidden_size)

    def forward(self, hidden_states):
        B, S, _ = hidden_states.shape
        residual = hidden_states
        x = self.pre_attention_layer_norm(hidden_states)
        use_fused = (self.fused and _native(x) and K.fused_attention_supported(S, self.head_dim))
        a = self._attention_fused(x, B, S) if use_fused else self._attention_unfused(x, B, S)
        a = self.attention_output_dropout(self.out_proj(a))
        h = a + residual

        shape = h.shape
        h2 = h.reshape(-1, shape[-1])
        residual = h2
        y = self.final_layer_norm(h2)
        y = self.ffn_2(self.activation(self.ffn_1(y)))
        y = self.final_dropout(y)
        return (residual + y).view(shape)


def load_opt_layer_state(block: TransformerBlock, opt_layer_state: dict) -> None:
    """Copies the weights of one HF OPT decoder layer (keys as in ``OPTDecoderLayer.state_dict()``)
    into the block — the in-repo version of ``_block_from_pretrained`` (model/transformer.py:110-125)."""
    m = {"self_attn_layer_norm": "pre_attention_layer_norm", "self_attn.k_proj": "k_proj",
         "self_attn.q_proj": "q_proj", "self_attn.v_proj": "v_proj", "self_attn.out_proj": "out_proj",
         "final_layer_norm": "final_layer_norm", "fc1": "ffn_1", "fc2": "ffn_2"}
    sd = {}
    for src, dst in m.items():
        for leaf in ("weight", "bias"):
            k = f"{src}.{leaf}"
            if k in opt_layer_state:
                sd[f"{dst}.{leaf}"] = opt_layer_state[k]
    block.load_state_dict(sd, strict=True)
