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
            Fn.mark_shared_grad(self.word_embeddings, 2)
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
            Fn.mark_shared_grad(self.word_embeddings, 2)   # survives deepcopy
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


# ------------------------------------------------------------------ HF parameter-name bridge
# The reference trains HF checkpoints (AutoModelForCausalLM.from_pretrained("bigscience/bloom-560m"
# / "facebook/opt-125m"), zero_dp_training.py:24).  These two functions map between HF's
# BloomForCausalLM / OPTForCausalLM state dicts and this module's fused layout (one QKV weight with
# columns Q | K | V, each heads x head_dim).
_LEAVES = ("weight", "bias")


def _bloom_qkv_from_hf(t: torch.Tensor, heads: int) -> torch.Tensor:
    """HF BLOOM stores query_key_value per head: rows ordered [head, (q, k, v), head_dim]."""
    d = t.shape[0] // (3 * heads)
    rest = t.shape[1:]
    return t.reshape(heads, 3, d, *rest).transpose(0, 1).reshape(3 * heads * d, *rest).contiguous()


def _bloom_qkv_to_hf(t: torch.Tensor, heads: int) -> torch.Tensor:
    d = t.shape[0] // (3 * heads)
    rest = t.shape[1:]
    return t.reshape(3, heads, d, *rest).transpose(0, 1).reshape(3 * heads * d, *rest).contiguous()


def causal_lm_to_hf_state_dict(model: "CausalLM") -> dict:
    """State dict with HF ``BloomForCausalLM`` / ``OPTForCausalLM`` key names."""
    sd, cfg, out = model.state_dict(), model.cfg, {}
    L, heads = cfg.num_hidden_layers, cfg.num_attention_heads
    if cfg.arch == "bloom":
        out["transformer.word_embeddings.weight"] = sd["word_embeddings"].clone()
        for leaf in _LEAVES:
            out[f"transformer.word_embeddings_layernorm.{leaf}"] = sd[f"embedding_ln.{leaf}"].clone()
            out[f"transformer.ln_f.{leaf}"] = sd[f"final_ln.{leaf}"].clone()
        names = {"input_ln": "input_layernorm", "attention.out": "self_attention.dense",
                 "post_attention_ln": "post_attention_layernorm", "mlp.fc1": "mlp.dense_h_to_4h",
                 "mlp.fc2": "mlp.dense_4h_to_h"}
        for i in range(L):
            for leaf in _LEAVES:
                out[f"transformer.h.{i}.self_attention.query_key_value.{leaf}"] = \
                    _bloom_qkv_to_hf(sd[f"layers.{i}.attention.qkv.{leaf}"], heads)
                for mine, hf in names.items():
                    out[f"transformer.h.{i}.{hf}.{leaf}"] = sd[f"layers.{i}.{mine}.{leaf}"].clone()
        out["lm_head.weight"] = (sd["word_embeddings"] if model.lm_head_weight is None else sd["lm_head_weight"]).clone()
        return out
    if cfg.arch == "opt":
        H = cfg.hidden_size
        p = "model.decoder."
        out[p + "embed_tokens.weight"] = sd["word_embeddings"].clone()
        out[p + "embed_positions.weight"] = sd["position_embeddings"].clone()
        for leaf in _LEAVES:
            out[p + f"final_layer_norm.{leaf}"] = sd[f"final_ln.{leaf}"].clone()
        names = {"input_ln": "self_attn_layer_norm", "attention.out": "self_attn.out_proj",
                 "post_attention_ln": "final_layer_norm", "mlp.fc1": "fc1", "mlp.fc2": "fc2"}
        for i in range(L):
            for leaf in _LEAVES:
                qkv = sd[f"layers.{i}.attention.qkv.{leaf}"]
                for j, n in enumerate(("q_proj", "k_proj", "v_proj")):
                    out[p + f"layers.{i}.self_attn.{n}.{leaf}"] = qkv[j * H:(j + 1) * H].clone()
                for mine, hf in names.items():
                    out[p + f"layers.{i}.{hf}.{leaf}"] = sd[f"layers.{i}.{mine}.{leaf}"].clone()
        out["lm_head.weight"] = (sd["word_embeddings"] if model.lm_head_weight is None else sd["lm_head_weight"]).clone()
        return out
    raise ValueError(f"no HF bridge for arch {cfg.arch!r}")


def causal_lm_from_hf_state_dict(model: "CausalLM", hf_sd: dict) -> "CausalLM":
    """Loads a HF ``BloomForCausalLM`` / ``OPTForCausalLM`` state dict (e.g. a downloaded
    bloom-560m / opt-125m checkpoint) into `model`, which must have been built from the matching
    config."""
    cfg, sd = model.cfg, {}
    L, heads = cfg.num_hidden_layers, cfg.num_attention_heads
    hf_sd = {k.removeprefix("module."): v for k, v in hf_sd.items()}
    if cfg.arch == "bloom":
        sd["word_embeddings"] = hf_sd["transformer.word_embeddings.weight"]
        for leaf in _LEAVES:
            sd[f"embedding_ln.{leaf}"] = hf_sd[f"transformer.word_embeddings_layernorm.{leaf}"]
            sd[f"final_ln.{leaf}"] = hf_sd[f"transformer.ln_f.{leaf}"]
        names = {"input_ln": "input_layernorm", "attention.out": "self_attention.dense",
                 "post_attention_ln": "post_attention_layernorm", "mlp.fc1": "mlp.dense_h_to_4h",
                 "mlp.fc2": "mlp.dense_4h_to_h"}
        for i in range(L):
            for leaf in _LEAVES:
                sd[f"layers.{i}.attention.qkv.{leaf}"] = _bloom_qkv_from_hf(
                    hf_sd[f"transformer.h.{i}.self_attention.query_key_value.{leaf}"], heads)
                for mine, hf in names.items():
                    sd[f"layers.{i}.{mine}.{leaf}"] = hf_sd[f"transformer.h.{i}.{hf}.{leaf}"]
    elif cfg.arch == "opt":
        p = "model.decoder."
        sd["word_embeddings"] = hf_sd[p + "embed_tokens.weight"]
        sd["position_embeddings"] = hf_sd[p + "embed_positions.weight"]
        for leaf in _LEAVES:
            sd[f"final_ln.{leaf}"] = hf_sd[p + f"final_layer_norm.{leaf}"]
        names = {"input_ln": "self_attn_layer_norm", "attention.out": "self_attn.out_proj",
                 "post_attention_ln": "final_layer_norm", "mlp.fc1": "fc1", "mlp.fc2": "fc2"}
        for i in range(L):
            for leaf in _LEAVES:
                sd[f"layers.{i}.attention.qkv.{leaf}"] = torch.cat(
                    [hf_sd[p + f"layers.{i}.self_attn.{n}.{leaf}"] for n in ("q_proj", "k_proj", "v_proj")], 0)
                for mine, hf in names.items():
                    sd[f"layers.{i}.{mine}.{leaf}"] = hf_sd[p + f"layers.{i}.{hf}.{leaf}"]
    else:
        raise ValueError(f"no HF bridge for arch {cfg.arch!r}")
    if model.lm_head_weight is not None:
        sd["lm_head_weight"] = hf_sd["lm_head.weight"]
    own = model.state_dict()
    missing = [k for k in own if k not in sd]
    if missing:
        raise KeyError(f"HF state dict lacks the sources of {missing[:4]}")
    with torch.no_grad():
        for k, v in own.items():
            if v.shape != sd[k].shape:
                raise ValueError(f"{k}: checkpoint shape {tuple(sd[k].shape)} != model shape {tuple(v.shape)}")
            v.copy_(sd[k].to(dtype=v.dtype, device=v.device))
    return model


def load_pretrained(model: nn.Module, path: str) -> nn.Module:
    """Loads a local HF checkpoint (a ``pytorch_model.bin`` / ``model.safetensors`` file, or a
    directory holding one) into a model built by ``build_model``: the offline stand-in for the
    reference's ``from_pretrained`` calls (data_parallel_training.py:30-31, zero_dp_training.py:24)."""
    import os
    if os.path.isdir(path):
        for fname in ("model.safetensors", "pytorch_model.bin"):
            if os.path.exists(os.path.join(path, fname)):
                path = os.path.join(path, fname)
                break
        else:
            raise FileNotFoundError(f"no model.safetensors / pytorch_model.bin under {path}")
    if path.endswith(".safetensors"):
        from safetensors.torch import load_file
        sd = load_file(path)
    else:
        sd = torch.load(path, map_location="cpu", weights_only=True)
    cfg = getattr(model, "cfg", None)
    if cfg is not None and cfg.arch == "bert":
        from .bert import from_hf_state_dict
        from_hf_state_dict(model, sd)
        return model
    return causal_lm_from_hf_state_dict(model, sd)
