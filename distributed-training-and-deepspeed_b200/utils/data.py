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


class WordHashTokenizer:
    """Offline tokenizer for real text when no HF tokenizer files are on the machine: lower-cased
    whitespace/punctuation split, each word hashed (FNV-1a) into ``[first_id, vocab_size)``.  Not
    a language model's vocabulary, but deterministic, and it gives real token statistics (Zipfian
    repeats, variable row lengths) to the kernels that care: embedding scatter-add, padding masks."""

    def __init__(self, vocab_size: int, model_max_length: int = 512, mask_token_id: int = 103,
                 pad_token_id: int = 0, first_id: int = 1000):
        self.vocab_size, self.model_max_length = vocab_size, model_max_length
        self.mask_token_id, self.pad_token_id = mask_token_id, pad_token_id
        self.first_id = min(first_id, vocab_size // 2)

    # Byte-level rules, identical to csrc/text_rows.cpp (the native packer used for whole files):
    # ASCII lower-casing; a word is a run of [a-z0-9_] or bytes >= 0x80; any other non-space byte
    # is a token of its own.
    def encode(self, text: str) -> list:
        out, span = [], self.vocab_size - self.first_id
        data = text.encode("utf-8")
        i, n = 0, len(data)

        def low(c):
            return c + 32 if 65 <= c <= 90 else c

        def is_word(c):
            return 97 <= c <= 122 or 48 <= c <= 57 or c == 95 or c >= 0x80

        while i < n:
            c = low(data[i])
            if c == 32 or 9 <= c <= 13:
                i += 1
                continue
            h = 0xcbf29ce484222325
            if is_word(c):
                while i < n and is_word(low(data[i])):
                    h = ((h ^ low(data[i])) * 0x100000001b3) & 0xFFFFFFFFFFFFFFFF
                    i += 1
            else:
                h = ((h ^ c) * 0x100000001b3) & 0xFFFFFFFFFFFFFFFF
                i += 1
            out.append(self.first_id + h % span)
        return out

    def encode_file(self, path: str, max_length: int, limit: int | None = None):
        """(ids [rows, max_length] int64 padded, lengths [rows]) for the non-empty lines of `path`
        through the multi-threaded native packer; None if the library is not built."""
        try:
            from ..ops import _lib
            L = _lib.lib()
        except Exception:  # noqa: BLE001
            return None
        import os
        cap = limit if limit is not None else max(1, sum(1 for _ in open(path, "rb")))
        ids = torch.empty(cap, max_length, dtype=torch.int64)
        lens = torch.empty(cap, dtype=torch.int32)
        rows = L.dtds_text_rows(os.fsencode(path), self.vocab_size, self.first_id, max_length, self.pad_token_id,
                                ids.data_ptr(), lens.data_ptr(), cap, 0)
        if rows < 0:
            raise OSError(f"dtds_text_rows({path}) failed with code {rows}")
        return ids[:rows], lens[:rows]


class TextFileDataset(Dataset):
    """Rows built the way the reference builds them from wikitext (util.py:40-53): one non-empty
    line per row, truncated / padded to ``max_length``, masking applied ONCE at build time (static
    masks); MLM rows follow the 15 % / 80-10-10 recipe of
    ``DataCollatorForLanguageModeling``, causal rows use ``labels = input_ids``; padding is -100
    in the labels.  `tokenizer` is a ``WordHashTokenizer`` or any object with ``encode(text)``
    (e.g. a HF tokenizer loaded from local files)."""

    def __init__(self, path: str, tokenizer, mlm: bool, max_length: int, mlm_probability: float = 0.15,
                 seed: int = 0, limit: int | None = None):
        g = torch.Generator().manual_seed(seed)
        pad = getattr(tokenizer, "pad_token_id", 0) or 0
        mask_id = getattr(tokenizer, "mask_token_id", 103)
        vocab = getattr(tokenizer, "vocab_size", None) or len(tokenizer)
        packed = tokenizer.encode_file(path, max_length, limit) if isinstance(tokenizer, WordHashTokenizer) else None
        if packed is not None:   # native path: whole file tokenised and padded in C++
            ids, lens = packed
            if ids.shape[0] == 0:
                raise ValueError(f"{path}: no non-empty lines")
            real = torch.arange(max_length)[None, :] < lens[:, None]
            self._finish(ids, real, mlm, mask_id, vocab, mlm_probability, g)
            return
        rows = []
        with open(path, encoding="utf-8") as f:
            for line in f:
                line = line.strip()
                if not line:
                    continue
                try:
                    toks = tokenizer.encode(line, add_special_tokens=True)
                except TypeError:
                    toks = tokenizer.encode(line)
                toks = list(toks)[:max_length]
                if toks:
                    rows.append(toks)
                if limit is not None and len(rows) >= limit:
                    break
        if not rows:
            raise ValueError(f"{path}: no non-empty lines")
        ids = torch.full((len(rows), max_length), pad, dtype=torch.int64)
        real = torch.zeros(len(rows), max_length, dtype=torch.bool)
        for r, toks in enumerate(rows):
            ids[r, :len(toks)] = torch.tensor(toks, dtype=torch.int64)
            real[r, :len(toks)] = True
        self._finish(ids, real, mlm, mask_id, vocab, mlm_probability, g)

    def _finish(self, ids, real, mlm, mask_id, vocab, mlm_probability, g):
        labels = ids.clone()
        if mlm:
            prob = torch.full(ids.shape, mlm_probability)
            prob[~real] = 0.0
            masked = torch.bernoulli(prob, generator=g).bool()
            labels[~masked] = -100
            replace = torch.bernoulli(torch.full(ids.shape, 0.8), generator=g).bool() & masked
            ids[replace] = mask_id
            rand = torch.bernoulli(torch.full(ids.shape, 0.5), generator=g).bool() & masked & ~replace
            ids[rand] = torch.randint(0, vocab, (int(rand.sum()),), generator=g)
        else:
            labels[~real] = -100
        self.input_ids, self.labels = ids, labels
        self.attention_mask = real

    def __len__(self):
        return self.input_ids.shape[0]

    def __getitem__(self, i):
        return {"input_ids": self.input_ids[i], "labels": self.labels[i]}

    def select(self, indices):
        idx = torch.as_tensor(list(indices), dtype=torch.long)
        if len(idx) and int(idx.max()) >= len(self):   # small files: wrap around like extra epochs
            idx = idx % len(self)
        out = object.__new__(TextFileDataset)
        out.input_ids, out.labels = self.input_ids[idx], self.labels[idx]
        out.attention_mask = self.attention_mask[idx]
        return out


def load_text_file(path: str, tokenizer, collator, max_length: int | None = None, seed: int = 0,
                   limit: int | None = None):
    """Real-text counterpart of `load_wikitext` for a local file with one passage per line (e.g.
    an extracted ``wiki.train.raw``)."""
    S = max_length or tokenizer.model_max_length
    tok = tokenizer if hasattr(tokenizer, "encode") else WordHashTokenizer(
        tokenizer.vocab_size, S, getattr(tokenizer, "mask_token_id", 103))
    return TextFileDataset(path, tok, collator.mlm, S, getattr(collator, "mlm_probability", 0.15), seed, limit)


def load_wikitext(tokenizer, collator, max_length: int | None = None, num_rows: int = 36718, seed: int = 0,
                  path: str | None = None):
    """Same signature as the reference (util.py:38); 36,718 is the row count of the wikitext-2
    train split.  Rows are generated lazily large enough for ``.select(range(n))``.  With `path`
    (or ``DTDS_TEXT_FILE``) the rows come from that local text file instead of the generator."""
    import os
    path = path or os.environ.get("DTDS_TEXT_FILE")
    if path:
        return load_text_file(path, tokenizer, collator, max_length, seed)
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
