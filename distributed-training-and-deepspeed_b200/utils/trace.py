"""Tracing helpers: NVTX ranges and device-side timers.

The reference has no profiler hooks at all (wall-clock totals with ``time.time()``
around the loops, data_parallel_training.py:47,62-63).  Here:

* ``nvtx_range(name)`` / ``annotate_model(model)`` emit NVTX ranges (per training step, per layer,
  per bucket collective, per optimizer step) when ``DTDS_NVTX=1`` — free when off, and visible in
  any NVTX-aware profiler timeline.
* ``DeviceTimer`` measures a region with CUDA events on the launching stream and reports the
  maximum over ranks, the only kind of time this repository quotes.
"""
from __future__ import annotations

import contextlib
import os

import torch

_ENABLED = os.environ.get("DTDS_NVTX", "0") not in ("0", "", "false", "False")


def nvtx_enabled() -> bool:
    return _ENABLED


def set_nvtx(enabled: bool) -> None:
    global _ENABLED
    _ENABLED = bool(enabled)


def _push(name: str) -> None:   # replaced by tests
    if torch.cuda.is_available():
        torch.cuda.nvtx.range_push(name)


def _pop() -> None:
    if torch.cuda.is_available():
        torch.cuda.nvtx.range_pop()


@contextlib.contextmanager
def nvtx_range(name: str):
    if not _ENABLED:
        yield
        return
    _push(name)
    try:
        yield
    finally:
        _pop()


def annotate_model(model: torch.nn.Module, depth: int = 2) -> list:
    """Wraps the forward of every sub-module down to `depth` levels in an NVTX range named after
    the module path (``encoders.3``, ``head``...).  Backward kernels show up under the autograd
    node names of the fused functions.  Returns the hook handles (``h.remove()`` to undo)."""
    handles = []
    for name, mod in model.named_modules():
        if not name or name.count(".") >= depth:
            continue
        label = f"{name} [{type(mod).__name__}]"

        def pre(_m, _args, label=label):
            if _ENABLED:
                _push(label)

        def post(_m, _args, _out):
            if _ENABLED:
                _pop()

        handles.append(mod.register_forward_pre_hook(pre))
        handles.append(mod.register_forward_hook(post, always_call=True))
    return handles


class DeviceTimer:
    """``with timer: ...`` accumulates device time of the enclosed launches (CUDA events on the
    current stream; host clock on CPU).  ``ms(reduce=True)`` synchronises, and returns the total —
    the maximum over ranks when a process group is up."""

    def __init__(self, device=None):
        self.device = torch.device(device) if device is not None else (
            torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu"))
        self._pairs: list = []
        self._host_ms = 0.0
        self.count = 0

    def __enter__(self):
        if self.device.type == "cuda":
            self._s = torch.cuda.Event(enable_timing=True)
            self._s.record(torch.cuda.current_stream(self.device))
        else:
            import time
            self._t0 = time.perf_counter()
        return self

    def __exit__(self, *exc):
        if self.device.type == "cuda":
            e = torch.cuda.Event(enable_timing=True)
            e.record(torch.cuda.current_stream(self.device))
            self._pairs.append((self._s, e))
        else:
            import time
            self._host_ms += (time.perf_counter() - self._t0) * 1e3
        self.count += 1
        return False

    def ms(self, reduce: bool = True) -> float:
        total = self._host_ms
        if self._pairs:
            torch.cuda.synchronize(self.device)
            total += sum(s.elapsed_time(e) for s, e in self._pairs)
            self._host_ms, self._pairs = total, []
        if reduce:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
                t = torch.tensor([total], dtype=torch.float64,
                                 device=self.device if dist.get_backend() == "nccl" else "cpu")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                return float(t.item())
        return total
