"""Layer-split model parallelism and GPipe pipeline parallelism, one process per stage.

The reference places contiguous module groups on several GPUs inside ONE process, hops
activations with ``tensor.to(device)`` (model/bert_mp.py:39-47,91-99) and wraps the same modules in
``torch.distributed.pipeline.sync.Pipe`` for GPipe (model/bert_mp.py:73-89).  Here every stage is
its own rank and the activation / gradient hand-off between neighbours is NCCL point-to-point
(``isend`` / ``irecv`` on NCCL's own streams), posted ahead of the compute that needs it so the
transfer of micro-batch k+1 overlaps the compute of micro-batch k.

Schedule (GPipe, synchronous): all forward micro-batches, then all backward micro-batches in
reverse order; gradients accumulate across micro-batches; ONE optimizer step per batch.
``chunks=1`` is the naive model-parallel case: exactly one stage is busy at any time, which is
what the idle-time table demonstrates.  ``checkpoint='except_last'`` recomputes the forward of all
micro-batches but the last during backward, like ``Pipe``'s default.

Device idle time is measured on the device: CUDA events bracket every forward/backward segment
of a stage; the gap between the end of one segment and the start of the next is idle time
(the reference stamps host ``time.time()`` in hooks without synchronising — quirk Q7).
"""
from __future__ import annotations

import time

import torch
import torch.distributed as dist

from ..ops import functional as F
from ..utils import trace as _trace


def split_sizes(n_items: int, n_groups: int) -> list[int]:
    """Group sizes of ``np.array_split(range(n_items), n_groups)``: the first ``n % k`` groups get
    one extra item (14 modules -> 2: 7/7, 3: 5/5/4, 4: 4/4/3/3, 8: 2,2,2,2,2,2,1,1)."""
    base, extra = divmod(n_items, n_groups)
    return [base + (1 if i < extra else 0) for i in range(n_groups)]


def split_ranges(n_items: int, n_groups: int) -> list[range]:
    out, start = [], 0
    for s in split_sizes(n_items, n_groups):
        out.append(range(start, start + s))
        start += s
    return out


class IdleTracker:
    """Per-device idle time: sum of gaps between consecutive busy segments."""

    def __init__(self, device: torch.device, verbose: bool = False, log=None):
        self.device = device
        self.cuda = device.type == "cuda"
        self.verbose = verbose
        self.log = log or (lambda m: None)
        self._segments = []  # (start, end, label): CUDA events or host timestamps
        self._open = None

    def enter(self, label: str):
        if self.cuda:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(torch.cuda.current_stream(self.device))
        else:
            ev = time.perf_counter()
        self._open = (ev, label)
        if _trace.nvtx_enabled():
            _trace._push(label)
        if self.verbose:
            self.log(f"Entering {label} on device {self._dev_index()}")

    def leave(self):
        if self._open is None:
            return
        start, label = self._open
        if self.cuda:
            end = torch.cuda.Event(enable_timing=True)
            end.record(torch.cuda.current_stream(self.device))
        else:
            end = time.perf_counter()
        self._segments.append((start, end, label))
        self._open = None
        if _trace.nvtx_enabled():
            _trace._pop()
        if self.verbose:
            self.log(f"Finished {label} on device {self._dev_index()}")

    def _dev_index(self):
        return self.device.index if self.cuda else 0

    def totals(self) -> tuple[float, int]:
        """(total idle milliseconds, number of gaps) over everything recorded so far."""
        if self.cuda:
            torch.cuda.synchronize(self.device)
        idle, n = 0.0, 0
        for (s0, e0, _), (s1, e1, _) in zip(self._segments, self._segments[1:]):
            gap = e0.elapsed_time(s1) if self.cuda else (s1 - e0) * 1e3
            idle += max(gap, 0.0)
            n += 1
        return idle, n

    def reset(self):
        self._segments.clear()
        self._open = None


class PipelineStage:
    """One pipeline stage living in this process."""

    def __init__(self, modules, stage: int, num_stages: int, device: torch.device, *, act_dtype,
                 group=None, chunks: int = 1, checkpoint: str = "except_last",
                 tracker: IdleTracker | None = None, graphs: bool | None = None, graph_warmup: int = 2,
                 extra_params=()):
        self.modules = list(modules)
        self.stage, self.num_stages = stage, num_stages
        self.device, self.act_dtype = device, act_dtype
        self.group = group
        self.chunks = chunks
        self.checkpoint = checkpoint
        self.tracker = tracker or IdleTracker(device)
        self.is_first, self.is_last = stage == 0, stage == num_stages - 1
        if graphs is None:
            import os
            graphs = os.environ.get("DTDS_PIPE_GRAPHS", "1") != "0"
        self.use_graphs, self.graph_warmup = bool(graphs), int(graph_warmup)
        self.extra_params = list(extra_params)   # parameters used by loss_fn (the head)
        self._graphs, self._calls = None, 0

    # ------------------------------------------------------------------------------------
    def _run(self, x):
        for m in self.modules:
            x = m(x)
        return x

    def _forward(self, x, mb: int, n_mb: int):
        recompute = self.checkpoint == "always" or (self.checkpoint == "except_last" and mb < n_mb - 1)
        if recompute and n_mb > 1 and torch.is_grad_enabled() and not self.is_first:
            return F.checkpoint(self._run, x)   # replays the fused-dropout seed stream in the recomputation
        return self._run(x)

    def _irecv(self, shape, src):
        buf = torch.empty(shape, dtype=self.act_dtype, device=self.device)
        return buf, dist.irecv(buf, src=src, group=self.group)

    def forward_backward(self, input_ids, labels, act_shape, loss_fn):
        """One GPipe batch.  `input_ids` is consumed by the first stage, `labels` by the last;
        `act_shape` = (batch, seq, hidden) of the full batch at stage boundaries;
        `loss_fn(stage_output, labels_mb, n_valid)` -> the micro-batch's share of the batch loss:
        the sum of its token losses divided by `n_valid`, the number of labelled tokens of the WHOLE
        batch (a 1-element float tensor on the device; nothing is read back to the host).  Returns
        the batch loss on the last stage, None elsewhere.

        With ``graphs=True`` (default on a GPU when every parameter of the stage writes its gradient
        in place into a flat buffer) the forward and the backward of every micro-batch slot are
        captured once as CUDA graphs after `graph_warmup` eager batches and replayed afterwards;
        only the NCCL hand-offs stay eager.  At 4 sequences per micro-batch a stage is otherwise
        bound by launching ~400 kernels per micro-batch from Python."""
        if self.use_graphs and self.device.type == "cuda":
            self._calls += 1
            key = (tuple(act_shape), tuple(input_ids.shape) if self.is_first else None)
            if self._graphs is None and self._calls > self.graph_warmup and self._graphable():
                self._capture(act_shape, input_ids, labels, loss_fn, key)
            if self._graphs is not None and self._graphs.key == key and torch.is_grad_enabled():
                return self._replay(input_ids, labels)
        return self._eager(input_ids, labels, act_shape, loss_fn)

    # ------------------------------------------------------------------------------------ graphs
    def _graphable(self) -> bool:
        ps = [p for m in self.modules for p in m.parameters() if p.requires_grad]
        ps += [p for p in self.extra_params if p.requires_grad]
        return all(getattr(p, "main_grad", None) is not None for p in ps)

    def _reset_grad_flags(self) -> None:
        """Gradient state as after optimizer.zero_grad(): the first writer of every flat-buffer
        gradient overwrites, later micro-batches accumulate (this gets baked into the graphs)."""
        seen = set()
        for m in self.modules:
            for p in m.parameters():
                seen.add(id(p))
                p._grad_written = False
                p._uses_pending = getattr(p, "_grad_uses", 1)
        for p in self.extra_params:
            if id(p) not in seen:
                p._grad_written = False
                p._uses_pending = getattr(p, "_grad_uses", 1)

    def _capture(self, act_shape, input_ids, labels, loss_fn, key) -> None:
        from types import SimpleNamespace
        from ..ops import kernels as K
        M, dev = self.chunks, self.device
        B = act_shape[0]
        mb_shape = (B // M,) + tuple(act_shape[1:])
        g = SimpleNamespace(key=key, pool=torch.cuda.graph_pool_handle(), stream=torch.cuda.Stream(dev),
                            seed=torch.zeros(1, dtype=torch.int64, device=dev), fwd=[None] * M, bwd=[None] * M,
                            y=[None] * M, loss=[None] * M, gx=[None] * M, x=[], gy=[], lab=[], n_valid=None)
        for _ in range(M):
            if self.is_first:
                g.x.append(torch.zeros((B // M,) + tuple(input_ids.shape[1:]), dtype=input_ids.dtype, device=dev))
            else:
                g.x.append(torch.zeros(mb_shape, dtype=self.act_dtype, device=dev).requires_grad_(True))
            if self.is_last:
                g.lab.append(torch.full((B // M,) + tuple(labels.shape[1:]), -100, dtype=labels.dtype, device=dev))
            else:
                g.gy.append(torch.zeros(mb_shape, dtype=self.act_dtype, device=dev))
        if self.is_last:
            g.n_valid = torch.ones(1, dtype=torch.float32, device=dev)
        torch.cuda.synchronize(dev)
        K.set_seed_offset_tensor(g.seed)   # fresh dropout masks per replay (bumped once per batch)
        try:
            for k in range(M):
                gr = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gr, pool=g.pool, stream=g.stream):
                    y = self._forward(g.x[k], k, M)
                    if self.is_last:
                        g.loss[k] = loss_fn(y, g.lab[k], g.n_valid)
                    else:
                        g.y[k] = y
                g.fwd[k] = gr
            self._reset_grad_flags()
            for k in reversed(range(M)):
                gr = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gr, pool=g.pool, stream=g.stream):
                    if self.is_last:
                        g.loss[k].backward()
                    else:
                        g.y[k].backward(g.gy[k])
                    if not self.is_first:
                        g.gx[k] = g.x[k].grad
                g.bwd[k] = gr
        finally:
            K.set_seed_offset_tensor(None)
        torch.cuda.synchronize(dev)
        seen, g.params = set(), []
        for p in [p for m in self.modules for p in m.parameters()] + list(self.extra_params):
            if p.requires_grad and id(p) not in seen:
                seen.add(id(p))
                g.params.append(p)
        self._graphs = g

    def _replay(self, input_ids, labels):
        g, M, tr = self._graphs, self.chunks, self.tracker
        prev_rank, next_rank = self.stage - 1, self.stage + 1
        sends = []
        with torch.no_grad():
            g.seed.add_(1)
            if self.is_first:
                for k, ids in enumerate(input_ids.chunk(M)):
                    g.x[k].copy_(ids, non_blocking=True)
            if self.is_last:
                lab = labels.to(self.device, non_blocking=True)
                g.n_valid.copy_((lab != -100).sum().clamp_min(1).to(torch.float32).reshape(1))
                for k, l in enumerate(lab.chunk(M)):
                    g.lab[k].copy_(l)
            # ---------------- forward
            pending = dist.irecv(g.x[0].detach(), src=prev_rank, group=self.group) if not self.is_first else None
            for k in range(M):
                if pending is not None:
                    pending.wait()
                    pending = dist.irecv(g.x[k + 1].detach(), src=prev_rank, group=self.group) if k + 1 < M else None
                tr.enter(f"forward pass (micro-batch {k})")
                g.fwd[k].replay()
                tr.leave()
                if not self.is_last:
                    sends.append(dist.isend(g.y[k].detach(), dst=next_rank, group=self.group))
            # ---------------- backward, reverse order
            pending = dist.irecv(g.gy[M - 1], src=next_rank, group=self.group) if not self.is_last else None
            for k in reversed(range(M)):
                if pending is not None:
                    pending.wait()
                    pending = dist.irecv(g.gy[k - 1], src=next_rank, group=self.group) if k > 0 else None
                tr.enter(f"backward pass (micro-batch {k})")
                g.bwd[k].replay()
                tr.leave()
                if not self.is_first:
                    sends.append(dist.isend(g.gx[k], dst=prev_rank, group=self.group))
            for w in sends:
                w.wait()
            # what the backward hooks would have told the optimizer: every gradient has been written
            # (FusedAdam.step zeroes the gradient of parameters that report no write)
            for p in g.params:
                p._grad_written = True
            if self.is_last:
                return torch.stack([l.detach() for l in g.loss]).sum()
        return None

    # ------------------------------------------------------------------------------------ eager
    def _eager(self, input_ids, labels, act_shape, loss_fn):
        M = self.chunks
        B = act_shape[0]
        assert B % M == 0, f"batch {B} is not divisible into {M} micro-batches"
        mb_shape = (B // M,) + tuple(act_shape[1:])
        prev_rank, next_rank = self.stage - 1, self.stage + 1
        tr = self.tracker
        inputs, outputs, losses, sends = [], [], [], []
        ids_mb = input_ids.chunk(M) if self.is_first else [None] * M
        lab_mb = labels.chunk(M) if self.is_last else [None] * M
        n_valid = None
        if self.is_last:   # stays on the device: the micro-batch losses are sum_k / n_valid
            n_valid = (labels.to(self.device, non_blocking=True) != -100).sum().clamp_min(1).to(torch.float32).reshape(1)
        # ---------------- forward: all micro-batches
        pending = self._irecv(mb_shape, prev_rank) if not self.is_first else None
        for k in range(M):
            if self.is_first:
                x = ids_mb[k].to(self.device, non_blocking=True)
            else:
                buf, work = pending
                work.wait()
                pending = self._irecv(mb_shape, prev_rank) if k + 1 < M else None  # overlaps compute
                x = buf.requires_grad_(True)
            tr.enter(f"forward pass (micro-batch {k})")
            y = self._forward(x, k, M)
            if self.is_last:
                losses.append(loss_fn(y, lab_mb[k].to(self.device, non_blocking=True), n_valid))
            tr.leave()
            inputs.append(x)
            if self.is_last:
                outputs.append(None)
            else:
                outputs.append(y)
                sends.append(dist.isend(y.detach(), dst=next_rank, group=self.group))
        # ---------------- backward: reverse order
        pending = self._irecv(mb_shape, next_rank) if not self.is_last else None
        for k in reversed(range(M)):
            if self.is_last:
                tr.enter(f"backward pass (micro-batch {k})")
                losses[k].backward()
                tr.leave()
            else:
                gbuf, work = pending
                work.wait()
                pending = self._irecv(mb_shape, next_rank) if k > 0 else None
                tr.enter(f"backward pass (micro-batch {k})")
                outputs[k].backward(gbuf)
                tr.leave()
            if not self.is_first:
                sends.append(dist.isend(inputs[k].grad, dst=prev_rank, group=self.group))
            outputs[k] = None
        for w in sends:
            w.wait()
        if self.is_last:
            return torch.stack([l.detach() for l in losses]).sum()
        return None
