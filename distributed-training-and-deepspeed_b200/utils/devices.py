"""Device probes and human-readable sizes (util.py:6-35 in the reference)."""
from __future__ import annotations

import math

import torch


def get_device() -> str:
    """"cuda" when a GPU is visible, else "cpu" (the reference also probes Apple "mps")."""
    if torch.cuda.is_available():
        return "cuda"
    if getattr(torch.backends, "mps", None) is not None and torch.backends.mps.is_available():
        return "mps"
    return "cpu"


def get_device_count() -> int:
    return torch.cuda.device_count() if torch.cuda.is_available() else 1


_UNITS = ("B", "KB", "MB", "GB", "TB", "PB", "EB", "ZB", "YB")


def format_size(size_bytes: float) -> str:
    """1024-based size with two decimals, e.g. 18253430784 -> "17.0 GB"."""
    if size_bytes == 0:
        return "0B"
    exp = min(int(math.floor(math.log(size_bytes, 1024))), len(_UNITS) - 1)
    return "%s %s" % (round(size_bytes / math.pow(1024, exp), 2), _UNITS[exp])
