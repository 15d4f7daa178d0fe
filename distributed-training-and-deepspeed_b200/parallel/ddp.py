"""Replicated data parallelism with bucketed gradient all-reduce overlapped with backward.

Plays the role of ``torch.nn.parallel.DistributedDataParallel`` + the c10d ``Reducer``
(data_parallel_training.py:44) with a B200-first design:

  * gradients live in ONE flat buffer in symmetric memory (parallel/flat.py); a bucket is a byte
    range of it, so nothing is copied into or out of buckets — the weight-gradient GEMMs write
    into the range the collective reduces in place;
  * buckets are formed over parameters in reverse registration order (≈ gradient-ready order)
    with a ``bucket_cap_mb`` MiB cap and a small first bucket, like the Reducer;
  * when the last gradient of a bucket has been enqueued, a two-shot P2P all-reduce kernel is
    launched on a side stream (event-ordered after the producing kernels): every rank pulls its
    1/n slice from all peers over NVLink, averages in fp32 (the ÷world and the bf16 cast are fused
    into the reduce), and pushes the result to all peers — overlapping the rest of backward;
  * the end of backward joins the side stream back into the compute stream (no host sync).

On CPU (gloo) or with ``DTDS_COMM=nccl`` the same bucket schedule runs on
``torch.distributed.all_reduce`` — that is the baseline the custom kernel is measured against.
"""
from __future__ import annotations

import contextlib
import os

import torch
import torch.distributed as dist
from torch import nn

from .. import comm
from ..utils.trace import nvtx_range
from .flat import FlatBuffers, grad_uses


class _Bucket:
    __slots__ = ("index", "lo", "hi", "params", "pending", "launched")

    def __init__(self, index):
        self.index, self.lo, self.hi = index, None, None
        self.params, self.pending, self.launched = [], 0, False


def assign_buckets(params, offsets, bucket_cap_mb: float, elem_size: int, first_bucket_mb: float = 1.0,
                   solo=()):
    """Reverse-order, size-capped bucket assignment.  Returns a list of (param_indices, lo, hi)
    with lo/hi element offsets into the flat buffer (contiguous because consecutive parameters
    are adjacent in the flat layout).  Parameter indices in `solo` get a bucket of their own."""
    cap = int(bucket_cap_mb * 1024 * 1024) // elem_size
    first = int(first_bucket_mb * 1024 * 1024) // elem_size
    out, cur, cur_elems = [], [], 0
    limit = min(first, cap) if first > 0 else cap
    # The parameters assigned last are the ones whose gradients arrive last: their buckets cannot
    # hide behind any compute.  Once no more than one cap's worth of ordinary parameters is left,
    # the cap drops to half of that remainder, so only a small bucket sits at the very end.
    remaining = sum(p.numel() for p in params if p.numel() <= cap)
    tail_cap = None
    for i in reversed(range(len(params))):
        n = params[i].numel()
        alone = i in solo
        if tail_cap is None and remaining <= cap and n <= cap:
            tail_cap = max(remaining // 2, min(first, cap) if first > 0 else 1)
        eff = limit if tail_cap is None else min(limit, tail_cap)
        if cur and (cur_elems + n > eff or alone):
            out.append(cur)
            cur, cur_elems, limit = [], 0, cap
        cur.append(i)
        cur_elems += n
        if n <= cap:
            remaining -= n
        if alone:
            out.append(cur)
            cur, cur_elems, limit = [], 0, cap
    if cur:
        out.append(cur)
    res = []
    for idxs in out:
        lo = min(offsets[i] for i in idxs)
        hi = max(offsets[i] + params[i].numel() for i in idxs)
        res.append((idxs, lo, hi))
    return res


class DistributedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, device_ids=None, bucket_cap_mb: float = 25,
                 process_group=None, grad_dtype: torch.dtype | None = None,
                 broadcast_parameters: bool = True, comm_blocks: int = 0):
        super().__init__()
        self.module = module
        self.group = process_group
        self.world = comm.get_world_size(process_group)
        self.rank = comm.get_rank(process_group)
        self.bucket_cap_mb = bucket_cap_mb
        self.comm_blocks = comm_blocks
        # blocks of the all-reduce kernels that run next to backward's GEMMs (default 32: enough
        # loads in flight for ~100 GB/s per direction while leaving the SMs to the GEMM CTAs)
        self._overlap_blocks = int(os.environ.get("DTDS_DDP_OVERLAP_BLOCKS", "0") or 0)
        params = [p for p in module.parameters() if p.requires_grad]
        p0 = params[0]
        self.device = p0.device
        if device_ids is not None and self.device.type == "cuda":
            assert self.device.index == torch.device("cuda", device_ids[0]).index
        nvls = os.environ.get("DTDS_NVLS", "auto").lower()
        want_mc = nvls in ("1", "force", "on") or (nvls == "auto" and self.world >= 4)
        self.flat = FlatBuffers(params, grad_dtype=grad_dtype, symmetric=True, group=process_group,
                                pad_total_to=64 * max(self.world, 1), multicast_grads=want_mc)
        self._nvls = getattr(self.flat.grad_buf, "mc_ptr", 0) != 0
        self._custom = self.flat.symmetric and comm.custom_kernels_active(self.flat.grad_flat)
        self._verify_replicas()
        if broadcast_parameters:
            self.flat.broadcast_params(0)
        # buckets
        esz = self.flat.grad_flat.element_size()
        self.buckets: list[_Bucket] = []
        self._bucket_of = {}
        for bi, (idxs, lo, hi) in enumerate(assign_buckets(self.flat.params, self.flat.offsets,
                                                          bucket_cap_mb, esz)):
            b = _Bucket(bi)
            # 16-byte aligned byte range (parameter starts are 64-element aligned)
            b.lo, b.hi = lo, (hi + 63) // 64 * 64
            b.params = [self.flat.params[i] for i in idxs]
            self.buckets.append(b)
            for p in b.params:
                self._bucket_of[id(p)] = b
        self.optimizer = None   # set by fuse_optimizer(): Adam per bucket right after its all-reduce
        self._require_sync = True
        self._accumulating = False
        self.expose_grads = True
        self._in_backward = False
        self._comm_stream = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None
        # per-bucket Adam (fuse_optimizer) runs on its own stream behind the bucket's all-reduce, so
        # the update of bucket k overlaps the all-reduce of bucket k+1
        self._opt_stream = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None
        self._opt_event = None
        self._events = []
        for i, p in enumerate(self.flat.params):
            p._on_grad_ready = self._param_ready
            p.register_post_accumulate_grad_hook(self._autograd_grad_hook)
        self._reset_iteration()

    # ------------------------------------------------------------------------------ setup
    def _verify_replicas(self):
        if self.world < 2:
            return
        sig = (len(self.flat.params), self.flat.total, tuple(tuple(p.shape) for p in self.flat.params[:8]))
        sigs = [None] * self.world
        dist.all_gather_object(sigs, sig, group=self.group)
        if any(s != sigs[0] for s in sigs):
            raise RuntimeError(f"DDP: parameter shapes differ across ranks: {sigs}")

    def _reset_iteration(self):
        if not self._accumulating:  # gradients of earlier no_sync() passes must be kept
            self.flat.reset_grad_state()
            for p in self.flat.params:
                p.grad = None
        else:
            for p in self.flat.params:
                p._uses_pending = grad_uses(p)
        self._accumulating = False
        for b in self.buckets:
            b.pending = len(b.params)
            b.launched = False
        self._events.clear()
        self._callback_queued = False

    # ------------------------------------------------------------------------------ hooks
    def _autograd_grad_hook(self, p):
        """Gradient arrived through ordinary autograd (module not built on the fused ops)."""
        if p.grad is None:
            return
        if p.grad.data_ptr() != p.main_grad.data_ptr():
            if p._grad_written:
                p.main_grad.add_(p.grad.to(p.main_grad.dtype))
            else:
                p.main_grad.copy_(p.grad)
        p.grad = None
        p._grad_written = True
        self._param_ready(p)

    def _param_ready(self, p):
        if not self._require_sync:
            return
        if not self._callback_queued:
            self._callback_queued = True
            torch.autograd.Variable._execution_engine.queue_callback(self._finalize_backward)
        b = self._bucket_of[id(p)]
        b.pending -= 1
        if b.pending == 0 and not b.launched:
            self._launch_bucket(b)

    def fuse_optimizer(self, optimizer) -> None:
        """Lets the engine run `optimizer.step_range` for a bucket on the communication stream as
        soon as that bucket's all-reduce has been enqueued: the (HBM-bound) Adam update overlaps the
        rest of backward instead of running after it (unpack fused with Adam).
        ``optimizer.step()`` afterwards only advances the step count."""
        self.optimizer = optimizer

    def _step_bucket(self, b, stream):
        if self.optimizer is None or self._accumulating:
            return
        self.optimizer.step_range(self.flat, b.lo, min(b.hi, self.flat.total), stream=stream)

    def _launch_bucket(self, b: _Bucket):
        with nvtx_range(f"ddp.bucket{b.index} all_reduce {(b.hi - b.lo) * self.flat.grad_flat.element_size() >> 10} KiB"):
            self._launch_bucket_impl(b)

    def _launch_bucket_impl(self, b: _Bucket):
        b.launched = True
        g = self.flat.grad_flat
        esz = g.element_size()
        nbytes = (b.hi - b.lo) * esz
        if self.world < 2:
            self._step_bucket(b, None)
            return
        if self._custom:
            cur = torch.cuda.current_stream(self.device)
            ready = torch.cuda.Event()
            ready.record(cur)
            self._comm_stream.wait_event(ready)
            self._wait_wgrad_stream(self._comm_stream, cur)
            ctx = comm.symm_context(self.group)
            # The last buckets of a step (the embedding tables and the first layer) become ready when
            # backward has almost no GEMMs left to hide them behind: they get the wide, bandwidth-
            # oriented launch (one block per SM, two vectors in flight per thread).  Earlier buckets
            # use the 64-register variant on 32 SMs' worth of blocks next to the resident GEMM CTAs.
            tail = b.index >= len(self.buckets) - 3
            fused_opt = self.optimizer is not None and not self._accumulating
            # Splitting a large tail bucket so that Adam of piece i runs under the all-reduce of piece
            # i+1 was measured (2 B200s): the wide all-reduce blocks and the Adam blocks do not fit on
            # an SM together, the pieces serialise and each pays the kernel's latency (5.39 vs 5.26
            # ms/step).  Kept behind DTDS_DDP_TAIL_CHUNKS for experiments; one piece by default.
            lo_hi = [(b.lo, b.hi)]
            if tail and fused_opt and nbytes >= (16 << 20) and os.environ.get("DTDS_DDP_TAIL_CHUNKS", "0") != "0":
                k = min(4, nbytes // (8 << 20))
                step = ((b.hi - b.lo) // k + 64 * self.world - 1) // (64 * self.world) * (64 * self.world)
                lo_hi = [(lo, min(lo + step, b.hi)) for lo in range(b.lo, b.hi, step)]
            for lo, hi in lo_hi:
                nb = (hi - lo) * esz
                with torch.cuda.stream(self._comm_stream):
                    s = torch.cuda.Event(enable_timing=comm.comms_logger.enabled)
                    e = torch.cuda.Event(enable_timing=comm.comms_logger.enabled)
                    s.record()
                    if self._nvls and nb >= (256 << 10) and g.dtype in (torch.bfloat16, torch.float32):
                        # in-switch reduction: multimem.ld_reduce + multimem.st (csrc/comm_p2p.cu)
                        from ..comm.symm import all_reduce_nvls
                        all_reduce_nvls(ctx, self.flat.grad_buf, lo * esz, nb, g.dtype,
                                        scale=1.0 / self.world,
                                        blocks=self.comm_blocks or (96 if tail else (self._overlap_blocks or 32)))
                    elif tail and not self.comm_blocks:
                        # wide, bandwidth-oriented launch: one block per SM, two vectors in flight per thread
                        ctx.all_reduce(self.flat.grad_buf, lo * esz, nb, g.dtype, scale=1.0 / self.world,
                                       algo=2, blocks=128)
                    else:
                        # 64-register variant on 32 SMs' worth of blocks next to the resident GEMM CTAs
                        ctx.all_reduce(self.flat.grad_buf, lo * esz, nb, g.dtype, scale=1.0 / self.world,
                                       algo=3, blocks=self.comm_blocks or self._overlap_blocks)
                    e.record()
                comm.comms_logger.record_events("all_reduce", nb, self.world, s, e)
                self._events.append(e)
                if fused_opt:
                    self._opt_stream.wait_event(e)
                    self.optimizer.step_range(self.flat, lo, min(hi, self.flat.total), stream=self._opt_stream)
                    self._opt_event = torch.cuda.Event()
                    self._opt_event.record(self._opt_stream)
        else:
            view = g[b.lo:b.hi]
            comm.all_reduce(view, group=self.group, backend="nccl", scale=1.0 / self.world)
            self._step_bucket(b, None)

    def _wait_wgrad_stream(self, waiter, cur):
        """Gradients of one bucket may come from the main stream and from the weight-gradient side
        stream (ops/functional.py); the collective must be ordered after both."""
        from ..ops.functional import wgrad_side_stream
        side = wgrad_side_stream(self.device)
        if side is not None and side != cur:
            waiter.wait_stream(side)

    def _finalize_backward(self):
        """End of backward: launch stragglers, join the comm stream, expose gradients."""
        if self.device.type == "cuda":
            cur = torch.cuda.current_stream(self.device)
            self._wait_wgrad_stream(cur, cur)
        if self._require_sync:
            for b in self.buckets:
                if not b.launched:  # parameters that got no gradient this step
                    for p in b.params:
                        if not p._grad_written:
                            p.main_grad.zero_()
                            p._grad_written = True
                    self._launch_bucket(b)
            if self._custom and self._events:
                cur = torch.cuda.current_stream(self.device)
                cur.wait_event(self._events[-1])
                if self._opt_event is not None:
                    cur.wait_event(self._opt_event)
                    self._opt_event = None
            if self.expose_grads and self.flat.grad_dtype == self.flat.param_dtype:
                for p in self.flat.params:  # torch-style optimizers read p.grad
                    p.grad = p.main_grad
        self._callback_queued = False

    # ------------------------------------------------------------------------------ API
    def forward(self, *args, **kwargs):
        if self.device.type == "cuda":  # DDP moves inputs to its device (reference quirk Q4)
            args = tuple(a.to(self.device, non_blocking=True) if torch.is_tensor(a) else a for a in args)
            kwargs = {k: (v.to(self.device, non_blocking=True) if torch.is_tensor(v) else v)
                      for k, v in kwargs.items()}
        if torch.is_grad_enabled():
            if self._require_sync:
                self._reset_iteration()
            else:
                if not self._accumulating:
                    self.flat.reset_grad_state()
                self._accumulating = True
                for p in self.flat.params:
                    p._uses_pending = grad_uses(p)
        return self.module(*args, **kwargs)

    @contextlib.contextmanager
    def no_sync(self):
        """Skip the all-reduce (gradient accumulation); gradients keep accumulating in place."""
        old = self._require_sync
        self._require_sync = False
        try:
            yield
        finally:
            self._require_sync = old

    def train(self, mode: bool = True):
        super().train(mode)
        return self

    def state_dict(self, *a, **k):
        return self.module.state_dict(*a, **k)

    def load_state_dict(self, *a, **k):
        return self.module.load_state_dict(*a, **k)
