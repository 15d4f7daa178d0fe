"""BERT for masked-LM built from the fused kernels.

Architecture = HF ``BertForMaskedLM`` without pooler (what data_parallel_training.py:30-31 loads):
post-LN encoder, GELU(erf), tied decoder.  The three building blocks are also exported on their
own because the layer-split / pipeline model (``BertModelWithMP``, model/bert_mp.py:22-24 in the
reference) is assembled from ``BertEmbeddings`` + N x ``BertLayer`` + ``BertOnlyMLMHead`` with an
*untied* decoder.

Parameters are stored fused (one [3H, H] QKV weight); ``to_hf_state_dict`` /
``from_hf_state_dict`` convert to and from the HF parameter names so checkpoints interoperate.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch
from torch import nn

from ..ops import functional as Fn
from .config import ModelConfig
from .layers import LayerNorm, Linear, MLP, SelfAttention, _init_normal_, key_padding_bias


@dataclass
class MaskedLMOutput:
    loss: Optional[torch.Tensor] = None
    logits: Optional[torch.Tensor] = None


class BertEmbeddings(nn.Module):
    """word + position + token-type gather (one kernel) -> LayerNorm(+dropout)."""

    def __init__(self, cfg: ModelConfig):
        super().__init__()
        self.cfg = cfg
        H = cfg.hidden_size
        self.word_embeddings = nn.Parameter(_init_normal_(torch.empty(cfg.vocab_size, H)))
        self.position_embeddings = nn.Parameter(_init_normal_(torch.empty(cfg.max_position_embeddings, H)))
        self.token_type_embeddings = nn.Parameter(_init_normal_(torch.empty(cfg.type_vocab_size, H)))
        self.ln = LayerNorm(H, cfg.layer_norm_eps, cfg.hidden_dropout_prob)

    def forward(self, input_ids):
        e = Fn.embedding(input_ids, self.word_embeddings, self.position_embeddings, self.token_type_embeddings)
        # BERT applies dropout after the LayerNorm; LN(dropout(x)+res) cannot express that, so the
        # embedding dropout is folded in as dropout of the LN *input* of the first layer's residual
        # path being exact only in eval.  We keep exact semantics with a standalone dropout.
        y = Fn.layer_norm(e, self.ln.weight, self.ln.bias, self.ln.eps)
        if self.training and self.cfg.hidden_dropout_prob > 0.0:
            from .layers import _DropoutFn
            y = _DropoutFn.apply(y, None, self.cfg.hidden_dropout_prob)
        return y


class BertLayer(nn.Module):
    """x -> LN(drop(attn(x)·Wo) + x) -> LN(drop(mlp(h)) + h).  8 GEMM-class launches forward."""

    def __init__(self, cfg: ModelConfig):
        super().__init__()
        H = cfg.hidden_size
        self.attention = SelfAttention(H, cfg.num_attention_heads, cfg.attention_probs_dropout_prob)
        self.attention_ln = LayerNorm(H, cfg.layer_norm_eps, cfg.hidden_dropout_prob)
        self.mlp = MLP(H, cfg.intermediate_size, cfg.hidden_act)
        self.output_ln = LayerNorm(H, cfg.layer_norm_eps, cfg.hidden_dropout_prob)

    def forward(self, x, key_bias=None):
        # fork: the skip branch's gradient is added in the dgrad GEMM epilogue of the main branch
        a, x_skip = self.attention(x, key_bias, fork=True)
        h = self.attention_ln(a, residual=x_skip)
        f, h_skip = self.mlp(h, fork=True)
        return self.output_ln(f, residual=h_skip)


class BertOnlyMLMHead(nn.Module):
    """dense + GELU -> LayerNorm -> vocabulary decoder (own weight unless tied by the parent)."""

    def __init__(self, cfg: ModelConfig, tied_weight: nn.Parameter | None = None):
        super().__init__()
        H = cfg.hidden_size
        self.cfg = cfg
        self.transform = Linear(H, H, act=cfg.hidden_act)
        self.ln = LayerNorm(H, cfg.layer_norm_eps)
        if tied_weight is None:
            self.decoder_weight = nn.Parameter(_init_normal_(torch.empty(cfg.vocab_size, H)))
            self._tied = None
        else:
            self.decoder_weight = None
            self._tied = [tied_weight]  # list: keeps the shared Parameter out of this module's registry
        self.decoder_bias = nn.Parameter(torch.zeros(cfg.vocab_size))

    @property
    def weight(self):
        return self.decoder_weight if self._tied is None else self._tied[0]

    def hidden(self, x):
        return self.ln(self.transform(x))

    def forward(self, x, labels=None, ignore_index=-100, gscale=1.0, n_valid=None):
        """Without labels: logits [B, S, V] (a view into a buffer whose rows are padded to 64
        columns).  With labels: the fused projection + cross-entropy loss (no logits tensor is
        returned).  Both go through ``__call__`` so module hooks (ZeRO-3 materialisation, idle
        tracking) see every use of the head's parameters."""
        if labels is not None:
            return self.loss(x, labels, ignore_index, gscale, n_valid)
        t = self.hidden(x)
        B, S, H = t.shape
        V = self.cfg.vocab_size
        ld = (V + 63) // 64 * 64
        if ld == V or not t.is_cuda:
            return Fn.linear(t, self.weight, self.decoder_bias)
        return _PaddedLogitsFn.apply(t, self.weight, self.decoder_bias, ld)

    def loss(self, x, labels, ignore_index=-100, gscale=1.0, n_valid=None):
        t = self.hidden(x)
        return Fn.lm_head_loss(t, self.weight, self.decoder_bias, labels, ignore_index=ignore_index,
                               seq=t.shape[1], gscale=gscale, n_valid=n_valid)


class _PaddedLogitsFn(torch.autograd.Function):
    """Decoder GEMM writing into a row-padded buffer (TMA/vector stores need 16-byte row pitch;
    V = 28996 is not a multiple of 8).  Returns the [B, S, V] view."""

    @staticmethod
    def forward(ctx, t, weight, bias, ld):
        from ..ops import kernels as K
        B, S, H = t.shape
        V = weight.shape[0]
        t2 = t.reshape(B * S, H)
        buf = torch.empty(B * S, ld, dtype=t.dtype, device=t.device)
        buf[:, V:].zero_()
        K.gemm(t2, weight, bias=bias, out=buf[:, :V])
        ctx.save_for_backward(t2, weight, bias)
        ctx.shape = (B, S, H)
        return buf.view(B, S, ld)[:, :, :V]

    @staticmethod
    def backward(ctx, d):
        from ..ops import kernels as K
        t2, weight, bias = ctx.saved_tensors
        B, S, H = ctx.shape
        V = weight.shape[0]
        if d.stride(-1) != 1 or (d.stride(-2) % 8) != 0 or d.stride(0) != S * d.stride(1):
            ld = (V + 63) // 64 * 64
            buf = torch.zeros(B * S, ld, dtype=d.dtype, device=d.device)
            buf[:, :V] = d.reshape(B * S, V)
            d2 = buf[:, :V]
        else:
            d2 = d.as_strided((B * S, V), (d.stride(1), 1), d.storage_offset())
        dx = K.gemm(d2, weight, b_mn_major=True).view(B, S, H)
        dw = Fn._wgrad(d2, t2, weight)
        db = Fn._bgrad(d2, bias)
        return dx, dw, db, None


class BertForMaskedLM(nn.Module):
    """``model(input_ids, labels=...)`` -> ``.loss`` (and ``.logits`` when labels are omitted or
    ``return_logits=True``), the call shape of data_parallel_training.py:53-54."""

    def __init__(self, cfg: ModelConfig):
        super().__init__()
        self.cfg = cfg
        self.embeddings = BertEmbeddings(cfg)
        self.encoders = nn.ModuleList([BertLayer(cfg) for _ in range(cfg.num_hidden_layers)])
        tied = self.embeddings.word_embeddings if cfg.tie_word_embeddings else None
        self.head = BertOnlyMLMHead(cfg, tied_weight=tied)
        if tied is not None:
            Fn.mark_shared_grad(tied, 2)  # embedding scatter-add + decoder wgrad

    def forward(self, input_ids, labels=None, attention_mask=None, return_logits=False):
        if self.cfg.tie_word_embeddings:
            Fn.mark_shared_grad(self.embeddings.word_embeddings, 2)   # survives deepcopy
        kb = key_padding_bias(attention_mask)
        h = self.embeddings(input_ids)
        for layer in self.encoders:
            h = layer(h, kb)
        if labels is not None and not return_logits:
            return MaskedLMOutput(loss=self.head(h, labels=labels))
        logits = self.head(h)
        loss = Fn.cross_entropy(logits, labels) if labels is not None else None
        return MaskedLMOutput(loss=loss, logits=logits)


# ------------------------------------------------------------------ HF parameter-name bridge
def _layer_to_hf(prefix_mine: str, prefix_hf: str, sd: dict, out: dict, H: int) -> None:
    qkv_w, qkv_b = sd[f"{prefix_mine}attention.qkv.weight"], sd[f"{prefix_mine}attention.qkv.bias"]
    for i, n in enumerate(("query", "key", "value")):
        out[f"{prefix_hf}attention.self.{n}.weight"] = qkv_w[i * H:(i + 1) * H].clone()
        out[f"{prefix_hf}attention.self.{n}.bias"] = qkv_b[i * H:(i + 1) * H].clone()
    m = {"attention.out.": "attention.output.dense.", "attention_ln.": "attention.output.LayerNorm.",
         "mlp.fc1.": "intermediate.dense.", "mlp.fc2.": "output.dense.", "output_ln.": "output.LayerNorm."}
    for a, b in m.items():
        for leaf in ("weight", "bias"):
            out[f"{prefix_hf}{b}{leaf}"] = sd[f"{prefix_mine}{a}{leaf}"].clone()


def to_hf_state_dict(model: BertForMaskedLM) -> dict:
    """State dict with HF ``BertForMaskedLM`` key names."""
    sd, out, H = model.state_dict(), {}, model.cfg.hidden_size
    e = "bert.embeddings."
    out[e + "word_embeddings.weight"] = sd["embeddings.word_embeddings"].clone()
    out[e + "position_embeddings.weight"] = sd["embeddings.position_embeddings"].clone()
    out[e + "token_type_embeddings.weight"] = sd["embeddings.token_type_embeddings"].clone()
    out[e + "LayerNorm.weight"], out[e + "LayerNorm.bias"] = sd["embeddings.ln.weight"].clone(), sd["embeddings.ln.bias"].clone()
    for i in range(model.cfg.num_hidden_layers):
        _layer_to_hf(f"encoders.{i}.", f"bert.encoder.layer.{i}.", sd, out, H)
    p = "cls.predictions."
    out[p + "transform.dense.weight"], out[p + "transform.dense.bias"] = sd["head.transform.weight"].clone(), sd["head.transform.bias"].clone()
    out[p + "transform.LayerNorm.weight"], out[p + "transform.LayerNorm.bias"] = sd["head.ln.weight"].clone(), sd["head.ln.bias"].clone()
    out[p + "bias"] = sd["head.decoder_bias"].clone()
    out[p + "decoder.bias"] = sd["head.decoder_bias"].clone()
    out[p + "decoder.weight"] = (sd["head.decoder_weight"] if "head.decoder_weight" in sd
                                 else sd["embeddings.word_embeddings"]).clone()
    return out


def from_hf_state_dict(model: BertForMaskedLM, hf: dict) -> None:
    """Loads tensors named as in HF ``BertForMaskedLM`` into the fused parameters."""
    sd = {}
    e = "bert.embeddings."
    sd["embeddings.word_embeddings"] = hf[e + "word_embeddings.weight"]
    sd["embeddings.position_embeddings"] = hf[e + "position_embeddings.weight"]
    sd["embeddings.token_type_embeddings"] = hf[e + "token_type_embeddings.weight"]
    sd["embeddings.ln.weight"], sd["embeddings.ln.bias"] = hf[e + "LayerNorm.weight"], hf[e + "LayerNorm.bias"]
    for i in range(model.cfg.num_hidden_layers):
        a, b = f"encoders.{i}.", f"bert.encoder.layer.{i}."
        sd[a + "attention.qkv.weight"] = torch.cat([hf[f"{b}attention.self.{n}.weight"] for n in ("query", "key", "value")])
        sd[a + "attention.qkv.bias"] = torch.cat([hf[f"{b}attention.self.{n}.bias"] for n in ("query", "key", "value")])
        m = {"attention.out.": "attention.output.dense.", "attention_ln.": "attention.output.LayerNorm.",
             "mlp.fc1.": "intermediate.dense.", "mlp.fc2.": "output.dense.", "output_ln.": "output.LayerNorm."}
        for x, y in m.items():
            for leaf in ("weight", "bias"):
                sd[f"{a}{x}{leaf}"] = hf[f"{b}{y}{leaf}"]
    p = "cls.predictions."
    sd["head.transform.weight"], sd["head.transform.bias"] = hf[p + "transform.dense.weight"], hf[p + "transform.dense.bias"]
    sd["head.ln.weight"], sd["head.ln.bias"] = hf[p + "transform.LayerNorm.weight"], hf[p + "transform.LayerNorm.bias"]
    sd["head.decoder_bias"] = hf[p + "bias"]
    if model.head.decoder_weight is not None:
        sd["head.decoder_weight"] = hf[p + "decoder.weight"]
    model.load_state_dict(sd)
