"""Communication event recorder.

Same data model the reference script consumes from DeepSpeed's CommsLogger
(zero_dp_training.py:101-107): ``comms_dict[op_name][msg_size] = [count, [latency_ms...],
[algbw...], [busbw...]]``, and the same summary table as ``dist.log_summary()``.  Differences, on
purpose: latencies are measured on the device with CUDA events on the issuing stream and are
resolved lazily at summary time, so profiling every collective (``prof_all``) does not insert a
host synchronisation per call (DeepSpeed 0.9.2 serialises comm and compute to time them).
"""
from __future__ import annotations

import contextlib
import time

import torch


def calc_bw(op_name: str, size_bytes: int, latency_ms: float, world: int):
    """(algbw, busbw) in Gbit/s, NCCL-tests conventions: all-reduce 2(n-1)/n, all-gather /
    reduce-scatter (n-1)/n on the full message, everything else factor 1."""
    if latency_ms <= 0:
        return 0.0, 0.0
    t = latency_ms / 1e3
    n = max(world, 1)
    if "all_reduce" in op_name or op_name == "inference_all_reduce":
        alg = size_bytes / t
        bus = alg * (2 * (n - 1) / n)
    elif "all_gather" in op_name or "reduce_scatter" in op_name or "zero_step" in op_name:
        alg = size_bytes / t
        bus = alg * ((n - 1) / n)
    else:
        alg = size_bytes / t
        bus = alg
    return alg * 8 / 1e9, bus * 8 / 1e9


def convert_size(size_bytes: int) -> str:
    if size_bytes == 0:
        return "0B"
    names = ("B", "KB", "MB", "GB", "TB")
    i, s = 0, float(size_bytes)
    while s >= 1024 and i < len(names) - 1:
        s /= 1024
        i += 1
    return f"{round(s, 2)} {names[i]}"


class CommsLogger:
    def __init__(self):
        self.comms_dict: dict = {}
        self.enabled = False
        self.verbose = False
        self.prof_all = True
        self.debug = False
        self.prof_ops: list = []
        self._pending: list = []  # (op, size, world, start_event, end_event)

    def configure(self, cfg: dict) -> None:
        self.enabled = bool(cfg.get("enabled", False))
        self.verbose = bool(cfg.get("verbose", False))
        self.prof_all = bool(cfg.get("prof_all", True))
        self.debug = bool(cfg.get("debug", False))
        self.prof_ops = list(cfg.get("prof_ops", []))

    def _wanted(self, op: str) -> bool:
        return self.enabled and (self.prof_all or op in self.prof_ops)

    @contextlib.contextmanager
    def record(self, op: str, size_bytes: int, world: int, tensor=None):
        if not self._wanted(op):
            yield
            return
        on_gpu = tensor is not None and tensor.is_cuda
        if on_gpu and torch.cuda.is_current_stream_capturing():
            yield   # timing events cannot be resolved from inside a CUDA graph; replays are not logged
            return
        if on_gpu:
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            yield
            e.record()
            self._pending.append((op, size_bytes, world, s, e))
        else:
            t0 = time.perf_counter()
            yield
            self.append_host(op, size_bytes, (time.perf_counter() - t0) * 1e3, world)

    def record_events(self, op: str, size_bytes: int, world: int, start, end) -> None:
        """For collectives launched on side streams by the engines (events already recorded)."""
        if self._wanted(op):
            self._pending.append((op, size_bytes, world, start, end))

    def append_host(self, op: str, size_bytes: int, latency_ms: float, world: int) -> None:
        if not self._wanted(op) and op != "log_summary_barrier":
            return
        self._append(op, size_bytes, latency_ms, world)

    def _append(self, op, size_bytes, latency_ms, world):
        alg, bus = calc_bw(op, size_bytes, latency_ms, world)
        by_size = self.comms_dict.setdefault(op, {})
        if size_bytes in by_size:
            rec = by_size[size_bytes]
            rec[0] += 1
            rec[1].append(latency_ms)
            rec[2].append(alg)
            rec[3].append(bus)
        else:
            by_size[size_bytes] = [1, [latency_ms], [alg], [bus]]
        if self.verbose:
            print(f"comm op: {op} | time (ms): {latency_ms:.3f} | msg size: {convert_size(size_bytes)} | "
                  f"algbw (Gbps): {alg:.2f} | busbw (Gbps): {bus:.2f}")

    def resolve(self) -> None:
        """Turns pending CUDA event pairs into latencies (one device synchronisation)."""
        if not self._pending:
            return
        torch.cuda.synchronize()
        pending, self._pending = self._pending, []
        for op, size, world, s, e in pending:
            self._append(op, size, s.elapsed_time(e), world)

    def total_latency_ms(self, skip=("log_summary_barrier",)) -> float:
        self.resolve()
        total = 0.0
        for op, by_size in self.comms_dict.items():
            if op in skip:
                continue
            for _, vals in sorted(by_size.items()):
                total += sum(vals[1])
        return total

    def log_all(self, print_log: bool = True) -> None:
        self.resolve()
        if not print_log:
            return
        hdr = f"{'Comm. Op': <20}{'Message Size': <20}{'Count': <20}{'Total Latency(ms)': <20}" \
              f"{'Avg Latency(ms)': <20}{'tput_avg (Gbps)': <20}{'busbw_avg (Gbps)': <20}"
        print(hdr)
        for op in self.comms_dict:
            print(op)
            for size, vals in sorted(self.comms_dict[op].items()):
                count, lats = vals[0], vals[1]
                total = sum(lats)
                avg = total / max(len(lats), 1)
                alg = sum(vals[2]) / max(len(vals[2]), 1)
                bus = sum(vals[3]) / max(len(vals[3]), 1)
                print(f"{' ': <20}{convert_size(size): <20}{count: <20}{total: <20.2f}{avg: <20.2f}"
                      f"{alg: <20.2f}{bus: <20.2f}")
