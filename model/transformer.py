"""``from model.transformer import TransformerBlock, MatMul`` — import path of the reference kept;
the implementation is dtds_b200.models.transformer_block.

Run as a script it checks the block against HF's OPT decoder layer: both get the SAME randomly
initialised weights (the reference downloads facebook/opt-125m, model/transformer.py:110-140,
which needs the network) and the maximum difference of their outputs is printed.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from dtds_b200.models.transformer_block import (MatMul, TransformerBlock,  # noqa: E402,F401
                                                load_opt_layer_state)


def _parity_check():
    import torch
    from transformers import OPTConfig
    from transformers.models.opt.modeling_opt import OPTDecoderLayer

    torch.manual_seed(0)
    cfg = OPTConfig(hidden_size=768, num_attention_heads=12, ffn_dim=3072, max_position_embeddings=2048,
                    dropout=0.0, do_layer_norm_before=True)
    try:
        hf = OPTDecoderLayer(cfg, layer_idx=0).eval()
    except TypeError:
        hf = OPTDecoderLayer(cfg).eval()
    block = TransformerBlock(cfg).eval()
    load_opt_layer_state(block, hf.state_dict())
    x = torch.randn(1, 64, cfg.hidden_size)
    with torch.no_grad():
        mine = block(x)
        theirs = hf(x)
        theirs = theirs[0] if isinstance(theirs, tuple) else theirs
    print(f"Model Outputs: {mine}")
    print(f"Reference-layer Outputs: {theirs}")
    print(f"max |difference| = {(mine - theirs).abs().max().item():.3e}")


if __name__ == "__main__":
    _parity_check()
