"""Allocator telemetry in the line format of ``deepspeed.runtime.utils.memory_status``
(called every step on rank 0 at zero_dp_training.py:93)."""
from __future__ import annotations

import torch

_prev = {"alloc": 0.0, "max_alloc": 0.0}


def memory_status(msg: str, print_rank: int = -1, reset_max: bool = False, rank: int | None = None) -> str:
    if rank is None:
        try:
            import torch.distributed as dist
            rank = dist.get_rank() if dist.is_initialized() else 0
        except Exception:  # noqa: BLE001
            rank = 0
    if print_rank != -1 and rank != print_rank:
        return ""
    if not torch.cuda.is_available():
        line = f"RANK={rank} MEMSTATS {msg} device=cpu (no CUDA allocator statistics)"
        print(line)
        return line
    torch.cuda.synchronize()
    gb = 1024 ** 3
    alloc = torch.cuda.memory_allocated() / gb
    max_alloc = torch.cuda.max_memory_allocated() / gb
    cache = torch.cuda.memory_reserved() / gb
    max_cache = torch.cuda.max_memory_reserved() / gb
    d_alloc, d_max = alloc - _prev["alloc"], max_alloc - _prev["max_alloc"]
    _prev["alloc"], _prev["max_alloc"] = alloc, max_alloc
    line = (f"RANK={rank} MEMSTATS {msg} device={torch.cuda.current_device()} "
            f"current alloc={alloc:0.4f}GB (delta={d_alloc:0.4f}GB max={max_alloc:0.4f}GB) "
            f"current cache={cache:0.4f}GB (delta={d_max:0.4f}GB max={max_cache:0.4f}GB)")
    print(line)
    if reset_max:
        torch.cuda.reset_peak_memory_stats()
    return line
