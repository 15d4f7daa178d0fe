"""Model families: BERT (MLM, tied or untied head), BLOOM/OPT-shaped causal LMs, the hookable
``TransformerBlock`` and the layer-split / pipeline ``BertModelWithMP``."""
from .config import ModelConfig, PRESETS, get_config, transformer_scale_config  # noqa: F401
from .bert import (BertEmbeddings, BertForMaskedLM, BertLayer, BertOnlyMLMHead,  # noqa: F401
                   from_hf_state_dict, to_hf_state_dict)
from .causal_lm import (CausalLM, DecoderLayer, build_model, causal_lm_from_hf_state_dict,  # noqa: F401
                        causal_lm_to_hf_state_dict, count_parameters, load_pretrained)
from .transformer_block import MatMul, TransformerBlock, load_opt_layer_state  # noqa: F401
