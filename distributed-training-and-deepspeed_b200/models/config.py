"""In-repo model configurations (no hub access is needed or attempted).

The presets reproduce the architectures the reference's scripts load by name:
``bert-base-cased`` / ``bert-large-cased`` (data_parallel_training.py:30, model_parallel_training.py:36),
``bigscience/bloom-560m`` (zero_dp_training.py:117) and ``facebook/opt-125m``
(scripts/launch-multinode.sh:5, model/transformer.py:111-113).  Parameter counts are asserted in
tests/test_models_cpu.py against SURVEY.md §2.6.
"""
from __future__ import annotations

import dataclasses
from dataclasses import dataclass


@dataclass
class ModelConfig:
    name: str = "custom"
    arch: str = "bert"                 # "bert" | "bloom" | "opt"
    vocab_size: int = 28996
    hidden_size: int = 768
    num_hidden_layers: int = 12
    num_attention_heads: int = 12
    intermediate_size: int = 3072
    max_position_embeddings: int = 512
    type_vocab_size: int = 2
    hidden_act: str = "gelu"           # gelu | gelu_tanh | relu
    hidden_dropout_prob: float = 0.1
    attention_probs_dropout_prob: float = 0.1
    layer_norm_eps: float = 1e-12
    tie_word_embeddings: bool = True
    position_offset: int = 0           # OPT stores positions with an offset of 2
    # OPTConfig-style aliases consumed by TransformerBlock (model/transformer.py:30-53)
    enable_bias: bool = True

    @property
    def head_dim(self) -> int:
        return self.hidden_size // self.num_attention_heads

    @property
    def ffn_dim(self) -> int:
        return self.intermediate_size

    @property
    def dropout(self) -> float:
        return self.hidden_dropout_prob

    def replace(self, **kw) -> "ModelConfig":
        return dataclasses.replace(self, **kw)


PRESETS = {
    "bert-base-cased": ModelConfig(name="bert-base-cased", arch="bert", vocab_size=28996, hidden_size=768,
                                   num_hidden_layers=12, num_attention_heads=12, intermediate_size=3072),
    "bert-large-cased": ModelConfig(name="bert-large-cased", arch="bert", vocab_size=28996, hidden_size=1024,
                                    num_hidden_layers=24, num_attention_heads=16, intermediate_size=4096),
    "bigscience/bloom-560m": ModelConfig(name="bigscience/bloom-560m", arch="bloom", vocab_size=250880,
                                         hidden_size=1024, num_hidden_layers=24, num_attention_heads=16,
                                         intermediate_size=4096, max_position_embeddings=2048,
                                         type_vocab_size=0, hidden_act="gelu_tanh", hidden_dropout_prob=0.0,
                                         attention_probs_dropout_prob=0.0, layer_norm_eps=1e-5),
    "facebook/opt-125m": ModelConfig(name="facebook/opt-125m", arch="opt", vocab_size=50272, hidden_size=768,
                                     num_hidden_layers=12, num_attention_heads=12, intermediate_size=3072,
                                     max_position_embeddings=2048, type_vocab_size=0, hidden_act="relu",
                                     hidden_dropout_prob=0.1, attention_probs_dropout_prob=0.0,
                                     layer_norm_eps=1e-5, position_offset=2),
    # small shapes for CPU tests and smoke runs (head_dim 64 so the fused attention applies)
    "bert-tiny": ModelConfig(name="bert-tiny", arch="bert", vocab_size=1000, hidden_size=128,
                             num_hidden_layers=2, num_attention_heads=2, intermediate_size=512,
                             max_position_embeddings=128),
    "bloom-tiny": ModelConfig(name="bloom-tiny", arch="bloom", vocab_size=1000, hidden_size=128,
                              num_hidden_layers=2, num_attention_heads=2, intermediate_size=512,
                              max_position_embeddings=128, type_vocab_size=0, hidden_act="gelu_tanh",
                              hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, layer_norm_eps=1e-5),
    "opt-tiny": ModelConfig(name="opt-tiny", arch="opt", vocab_size=1000, hidden_size=128,
                            num_hidden_layers=2, num_attention_heads=2, intermediate_size=512,
                            max_position_embeddings=128, type_vocab_size=0, hidden_act="relu",
                            hidden_dropout_prob=0.1, attention_probs_dropout_prob=0.0, layer_norm_eps=1e-5,
                            position_offset=2),
}
ALIASES = {"base": "bert-base-cased", "large": "bert-large-cased", "bloom-560m": "bigscience/bloom-560m",
           "opt-125m": "facebook/opt-125m"}


def get_config(name: str) -> ModelConfig:
    key = ALIASES.get(name, name)
    if key not in PRESETS:
        raise KeyError(f"unknown model '{name}'; known: {sorted(PRESETS)}")
    return PRESETS[key].replace()


def transformer_scale_config(hidden_size: int, num_layers: int, heads: int | None = None,
                             vocab_size: int = 50272, seq: int = 2048, arch: str = "opt") -> ModelConfig:
    """A multi-billion-parameter decoder sized by hand (BASELINE.json config 5)."""
    heads = heads or hidden_size // 64
    return ModelConfig(name=f"{arch}-h{hidden_size}-l{num_layers}", arch=arch, vocab_size=vocab_size,
                       hidden_size=hidden_size, num_hidden_layers=num_layers, num_attention_heads=heads,
                       intermediate_size=4 * hidden_size, max_position_embeddings=seq, type_vocab_size=0,
                       hidden_act="relu" if arch == "opt" else "gelu_tanh", hidden_dropout_prob=0.0,
                       attention_probs_dropout_prob=0.0, layer_norm_eps=1e-5,
                       position_offset=2 if arch == "opt" else 0)
