"""Offline stand-ins for what the reference downloads from the Hugging Face hub at start-up.

The reference calls `BertForMaskedLM.from_pretrained("bert-base-cased")`,
`AutoTokenizer.from_pretrained("bert-base-cased")` (data_parallel_training.py:30-32 of the reference)
and `datasets.load_dataset("wikitext", "wikitext-2-v1")` (util.py:47).  Both libraries resolve a name
to a LOCAL DIRECTORY of that name before they try the hub, so the reference runs unmodified from a
working directory that contains:

    bert-base-cased/   config.json (the published bert-base-cased configuration), model.safetensors
                       (random-init weights of that architecture), vocab.txt (28,996 synthetic
                       WordPiece entries, special tokens at the published ids), tokenizer_config.json
    wikitext/          README.md (declares the `wikitext-2-v1` config) + train.jsonl: synthetic
                       passages, every one longer than 512 word pieces, so that after the reference's
                       `padding='max_length', truncation=True` every row is 512 real tokens (the
                       throughput of a BERT step does not depend on the token values)

Weights are random because there is no network; the arithmetic per step is the same.
"""
from __future__ import annotations

import json
import os
import random

BERT = {
    "bert-base-cased": dict(hidden_size=768, num_hidden_layers=12, num_attention_heads=12, intermediate_size=3072),
    "bert-large-cased": dict(hidden_size=1024, num_hidden_layers=24, num_attention_heads=16, intermediate_size=4096),
}
VOCAB = 28996


def _words(n: int):
    """Deterministic pronounceable pseudo-words (no network, no word list on the box)."""
    rnd = random.Random(7)
    cons, vow = "bcdfghjklmnprstvwz", "aeiou"
    seen, out = set(), []
    while len(out) < n:
        w = "".join(rnd.choice(cons) + rnd.choice(vow) for _ in range(rnd.randint(2, 4)))
        if rnd.random() < 0.15:
            w = w.capitalize()
        if w not in seen:
            seen.add(w)
            out.append(w)
    return out


def _base_tokens():
    toks = ["[PAD]"] + [f"[unused{i}]" for i in range(1, 100)] + ["[UNK]", "[CLS]", "[SEP]", "[MASK]"]
    toks += [f"[unused{i}]" for i in range(100, 102)]
    toks += list("!\"#$%&'()*+,-./0123456789:;<=>?@ABCDEFGHIJKLMNOPQRSTUVWXYZ[\\]^_`abcdefghijklmnopqrstuvwxyz{|}~")
    toks += [f"##{c}" for c in "abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ0123456789"]
    return toks


def make_vocab(path: str):
    toks = _base_tokens()
    toks += _words(VOCAB - len(toks))
    assert len(toks) == VOCAB and len(set(toks)) == VOCAB
    with open(path, "w") as f:
        f.write("\n".join(toks) + "\n")


def make_bert(workdir: str, name: str):
    d = os.path.join(workdir, name)
    if os.path.isfile(os.path.join(d, "model.safetensors")) and os.path.isfile(os.path.join(d, "vocab.txt")):
        return d
    import torch
    from transformers import BertConfig, BertForMaskedLM
    os.makedirs(d, exist_ok=True)
    cfg = BertConfig(vocab_size=VOCAB, max_position_embeddings=512, type_vocab_size=2, hidden_act="gelu",
                     hidden_dropout_prob=0.1, attention_probs_dropout_prob=0.1, layer_norm_eps=1e-12,
                     architectures=["BertForMaskedLM"], **BERT[name])
    torch.manual_seed(0)
    BertForMaskedLM(cfg).save_pretrained(d)
    make_vocab(os.path.join(d, "vocab.txt"))
    with open(os.path.join(d, "tokenizer_config.json"), "w") as f:
        json.dump({"do_lower_case": False, "model_max_length": 512, "tokenizer_class": "BertTokenizer",
                   "unk_token": "[UNK]", "sep_token": "[SEP]", "pad_token": "[PAD]", "cls_token": "[CLS]",
                   "mask_token": "[MASK]"}, f)
    return d


def make_wikitext(workdir: str, rows: int):
    d = os.path.join(workdir, "wikitext")
    marker = os.path.join(d, f".rows{rows}")
    if os.path.isfile(marker):
        return d
    os.makedirs(d, exist_ok=True)
    words = _words(VOCAB - len(_base_tokens()))
    rnd = random.Random(11)
    with open(os.path.join(d, "train.jsonl"), "w") as f:
        for _ in range(rows):
            f.write(json.dumps({"text": " " + " ".join(rnd.choices(words, k=560)) + " \n"}) + "\n")
    with open(os.path.join(d, "README.md"), "w") as f:
        f.write("---\nconfigs:\n- config_name: wikitext-2-v1\n  data_files:\n  - split: train\n    path: train.jsonl\n---\n"
                "Synthetic wikitext-shaped passages for the offline reference arm (baseline/make_assets.py).\n")
    for old in os.listdir(d):
        if old.startswith(".rows"):
            os.remove(os.path.join(d, old))
    open(marker, "w").close()
    return d


def make_all(workdir: str, model: str = "bert-base-cased", rows: int = 1024):
    os.makedirs(workdir, exist_ok=True)
    make_bert(workdir, model)
    make_wikitext(workdir, rows)
    return workdir


if __name__ == "__main__":
    import sys
    print(make_all(sys.argv[1] if len(sys.argv) > 1 else "/tmp/dtds_ref_work"))
