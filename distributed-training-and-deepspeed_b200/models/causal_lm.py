"""Decoder-only causal LMs in the shapes the ZeRO script trains (zero_dp_training.py:24,117;
scripts/launch-multinode.sh:5): BLOOM (ALiBi, embedding LayerNorm, GELU-tanh) and OPT (learned
positions with offset 2, ReLU).  Both are pre-LN with a final LayerNorm and a tied LM head; the
loss uses the causal-LM convention (labels = input_ids, shifted inside the loss kernel).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch
from torch import nn

from ..ops import functional as Fn
from .config import ModelConfig
from .layers import Dropout, LayerNorm, MLP, SelfAttention, _init_normal_, key_padding_bias


@dataclass
class CausalLMOutput:
    loss: Optional[torch.Tensor] = None
    logits: Optional[torch.Tensor] = None


class DecoderLayer(nn.Module):
    """x + drop(attn(LN(x))) ; h + drop(mlp(LN(h))).  With dropout 0 (BLOOM) both residual adds
    run in the epilogue of the producing GEMM."""

    def __init__(self, cfg: ModelConfig):
        super().__init__()
        H = cfg.hidden_size
        self.input_ln = LayerNorm(H, cfg.layer_norm_eps)
        self.attention = SelfAttention(H, cfg.num_attention_heads, cfg.attention_probs_dropout_prob,
                                       causal=True, alibi=(cfg.arch == "bloom"))
        self.post_attention_ln = LayerNorm(H, cfg.layer_norm_eps)
        self.mlp = MLP(H, cfg.intermediate_size, cfg.hidden_act)
        self.drop = Dropout(cfg.hidden_dropout_prob)

    def forward(self, x, key_bias=None):
        fuse_res = not (self.training and self.drop.p > 0.0)
        # fork: the skip branch's gradient is added inside the LayerNorm backward kernel
        n, x_skip = self.input_ln(x, fork=True)
        a = self.attention(n, key_bias, residual=x_skip if fuse_res else None)
        h = a if fuse_res else self.drop(a, residual=x_skip)
        n2, h_skip = self.post_attention_ln(h, fork=True)
        f = self.mlp(n2, residual=h_skip if fuse_res else None)
        return f if fuse_res else self.drop(f, residual=h_skip)


class CausalLM(nn.Module):
    """``model(input_ids, labels=input_ids)`` -> ``.loss`` as in zero_dp_training.py:85-86."""

    def __init__(self, cfg: ModelConfig):
        super().__init__()
        self.cfg = cfg
        H = cfg.hidden_size
        self.word_embeddings = nn.Parameter(_init_normal_(torch.empty(cfg.vocab_size, H)))
        if cfg.arch == "opt":
            self.position_embeddings = nn.Parameter(
                _init_normal_(torch.empty(cfg.max_position_embeddings + cfg.position_offset, H)))
            self.embedding_ln = None
        else:
            self.position_embeddings = None
            self.embedding_ln = LayerNorm(H, cfg.layer_norm_eps)
        self.layers = nn.ModuleList([DecoderLayer(cfg) for _ in range(cfg.num_hidden_layers)])
        self.final_ln = LayerNorm(H, cfg.layer_norm_eps)
        if cfg.tie_word_embeddings:
            self.lm_head_weight = None
            Fn.mark_shared_grad(self.word_embeddings, 2, sparse_rows=True)
        else:
            self.lm_head_weight = nn.Parameter(_init_normal_(torch.empty(cfg.vocab_size, H)))

    @property
    def head_weight(self):
        return self.word_embeddings if self.lm_head_weight is None else self.lm_head_weight

    def embed(self, input_ids):
        h = Fn.embedding(input_ids, self.word_embeddings, self.position_embeddings, None,
                         self.cfg.position_offset)
        if self.embedding_ln is not None:
            h = self.embedding_ln(h)
        return h

    def forward(self, input_ids, labels=None, attention_mask=None, return_logits=False):
        if self.lm_head_weight is None:
            Fn.mark_shared_grad(self.word_embeddings, 2, sparse_rows=True)   # survives deepcopy
        kb = key_padding_bias(attention_mask)
        h = self.embed(input_ids)
        for layer in self.layers:
            h = layer(h, kb)
        h = self.final_ln(h)
        if labels is not None and not return_logits:
            loss = Fn.lm_head_loss(h, self.head_weight, None, labels, shift=True, seq=h.shape[1])
            return CausalLMOutput(loss=loss)
        logits = Fn.linear(h, self.head_weight, None)
        loss = Fn.cross_entropy(logits, labels, shift=True) if labels is not None else None
        return CausalLMOutput(loss=loss, logits=logits)


def build_model(cfg: ModelConfig) -> nn.Module:
    """Factory used by the CLIs: BERT-style MLM or decoder-only causal LM by ``cfg.arch``."""
    if cfg.arch == "bert":
        from .bert import BertForMaskedLM
        return BertForMaskedLM(cfg)
    return CausalLM(cfg)


def count_parameters(model: nn.Module) -> int:
    seen, n = set(), 0
    for p in model.parameters():
        if id(p) not in seen:
            seen.add(id(p))
            n += p.numel()
    return n
