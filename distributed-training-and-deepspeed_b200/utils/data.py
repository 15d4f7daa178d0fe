"""Synthetic token datasets shaped like the reference's ``load_wikitext`` output (util.py:38-60).

The reference tokenises wikitext-2 with the model's tokenizer, pads every row to the maximum
length (512) and either applies a *static*, one-time MLM masking (15 % of the non-special
positions become labels; 80 % of those are replaced by [MASK], 10 % by a random token, 10 % kept)
or copies ``input_ids`` to ``labels`` for causal LMs.  There is no network here, so rows are drawn
from a seeded generator with the same shapes, dtypes and label conventions (-100 = ignored).
``max_length=None`` means the tokenizer's ``model_max_length`` (512), as in the reference.
"""
from __future__ import annotations

import torch
from torch.utils.data import Dataset


class SyntheticTextDataset(Dataset):
    """Rows ``{"input_ids": int64[S], "labels": int64[S]}``; supports ``.select(range(n))`` like a
    HF dataset (data_parallel_training.py:36)."""

    def __init__(self, num_rows: int, seq_len: int, vocab_size: int, mlm: bool, seed: int = 0,
                 mask_token_id: int = 103, mlm_probability: float = 0.15, special_below: int = 1000,
                 pad_fraction: float = 0.0):
        g = torch.Generator().manual_seed(seed)
        lo = min(special_below, vocab_size // 2)
        ids = torch.randint(lo, vocab_size, (num_rows, seq_len), generator=g, dtype=torch.int64)
        special = torch.zeros(num_rows, seq_len, dtype=torch.bool)
        if pad_fraction > 0:  # wikitext rows are short: most of a 512-token row is padding (id 0)
            lengths = (torch.rand(num_rows, generator=g) * (1 - pad_fraction) * seq_len).long().clamp(min=8)
            special = torch.arange(seq_len)[None, :] >= lengths[:, None]
            ids[special] = 0
        if mlm:
            labels = ids.clone()
            prob = torch.full(ids.shape, mlm_probability)
            prob[special] = 0.0
            masked = torch.bernoulli(prob, generator=g).bool()
            labels[~masked] = -100
            replace = torch.bernoulli(torch.full(ids.shape, 0.8), generator=g).bool() & masked
            ids[replace] = mask_token_id
            rand = torch.bernoulli(torch.full(ids.shape, 0.5), generator=g).bool() & masked & ~replace
            ids[rand] = torch.randint(lo, vocab_size, (int(rand.sum()),), generator=g)
        else:
            labels = ids.clone()
        self.input_ids, self.labels = ids, labels

    def __len__(self):
        return self.input_ids.shape[0]

    def __getitem__(self, i):
        return {"input_ids": self.input_ids[i], "labels": self.labels[i]}

    def select(self, indices):
        idx = torch.as_tensor(list(indices), dtype=torch.long)
        out = object.__new__(SyntheticTextDataset)
        out.input_ids, out.labels = self.input_ids[idx], self.labels[idx]
        return out


class _SyntheticTokenizer:
    """Stand-in for ``AutoTokenizer``: carries vocabulary size and maximum length."""

    def __init__(self, vocab_size: int, model_max_length: int = 512, mask_token_id: int = 103):
        self.vocab_size, self.model_max_length, self.mask_token_id = vocab_size, model_max_length, mask_token_id


class _SyntheticCollator:
    """Stand-in for ``DataCollatorForLanguageModeling`` (only the ``mlm`` switch matters)."""

    def __init__(self, tokenizer, mlm: bool = True, mlm_probability: float = 0.15):
        self.tokenizer, self.mlm, self.mlm_probability = tokenizer, mlm, mlm_probability


def synthetic_tokenizer(cfg) -> _SyntheticTokenizer:
    return _SyntheticTokenizer(cfg.vocab_size, min(512, cfg.max_position_embeddings))


def synthetic_collator(tokenizer, mlm: bool = True) -> _SyntheticCollator:
    return _SyntheticCollator(tokenizer, mlm)


def load_wikitext(tokenizer, collator, max_length: int | None = None, num_rows: int = 36718, seed: int = 0):
    """Same signature as the reference (util.py:38); 36,718 is the row count of the wikitext-2
    train split.  Rows are generated lazily large enough for ``.select(range(n))``."""
    S = max_length or tokenizer.model_max_length
    return LazySyntheticDataset(num_rows, dict(seq_len=S, vocab_size=tokenizer.vocab_size, mlm=collator.mlm,
                                               seed=seed, mask_token_id=tokenizer.mask_token_id,
                                               mlm_probability=getattr(collator, "mlm_probability", 0.15)))


class LazySyntheticDataset(Dataset):
    """Defers row generation until ``select`` (or first access) so asking for the 36,718-row split
    and then keeping ``batch_size * steps`` rows — what every script does — costs only those rows."""

    def __init__(self, num_rows: int, kwargs: dict):
        self.num_rows, self.kwargs, self._real = num_rows, kwargs, None

    def _materialize(self):
        if self._real is None:
            self._real = SyntheticTextDataset(self.num_rows, **self.kwargs)
        return self._real

    def select(self, indices):
        n = len(list(indices))
        assert n <= self.num_rows, f"asked for {n} rows, the split has {self.num_rows}"
        return SyntheticTextDataset(n, **self.kwargs)

    def __len__(self):
        return self.num_rows

    def __getitem__(self, i):
        return self._materialize()[i]
