"""ZeRO data parallelism (stages 0-3) behind a DeepSpeed-shaped API.

``initialize(model=..., model_parameters=..., config=ds_config)`` returns
``(engine, optimizer, None, None)`` like ``deepspeed.initialize`` (zero_dp_training.py:46-50) and
the engine exposes what the reference script calls: ``engine(...)``, ``engine.backward(loss)``,
``engine.step()``, ``engine.train()``, ``engine.zero_optimization_stage()``, and
``optimizer.param_groups[0]`` with ``lr / betas / eps / params`` (zero_dp_training.py:52-57,85-90).
Config keys honoured: ``train_micro_batch_size_per_gpu``, ``optimizer.type/params``,
``comms_logger.*``, ``zero_optimization.stage / reduce_bucket_size / allgather_bucket_size /
overlap_comm``, ``bf16.enabled``.

B200-first design:
  * parameters and gradients live in flat symmetric-memory buffers (parallel/flat.py); a bucket is
    a contiguous range of whole parameters in reverse registration order, capped at
    ``reduce_bucket_size`` elements; EVERY bucket is split evenly over the ranks, so all ranks
    reduce, update and publish 1/n of each bucket concurrently over NVSwitch (DeepSpeed 0.9.2
    instead issues one ``dist.reduce`` per (owner, slice) on a ring-era topology);
  * stages 1-3 run ONE kernel per bucket (csrc/comm_p2p.cu: zero_rs_adam_ag_kernel): pull the
    owned gradient slice from all peers (reduce-scatter), Adam on the fp32 master partition,
    push the updated bf16 parameters to all peers (all-gather) — no flatten/unflatten copies, no
    separate optimizer launch, no separate all-gather;
  * stage 2 launches that kernel from gradient-ready hooks on a side stream while backward is
    still running; stage 1 launches it for every bucket at ``step()``; stage 0 all-reduces the
    whole gradient buffer after backward and runs the flat fused Adam;
  * stage 3 additionally keeps only the owned 1/n shard of every parameter resident between
    uses (see ``Zero3ParamManager``).
On CPU/gloo, or with ``DTDS_COMM=nccl``, the same schedule runs on library collectives
(reduce_scatter_tensor / all_gather_into_tensor + the flat Adam kernel): that is the baseline.
"""
from __future__ import annotations

import math

import os

import torch
from torch import nn

from .. import comm
from ..ops import kernels as K
from ..utils.trace import nvtx_range
from .flat import FlatBuffers


class _ZBucket:
    __slots__ = ("index", "lo", "hi", "per", "params", "pending", "launched", "state_off", "goff", "prev_user",
                 "done_event")

    def __init__(self, index, lo, hi, world, params):
        self.index, self.lo, self.hi = index, lo, hi
        assert (hi - lo) % (8 * world) == 0, \
            f"bucket [{lo}, {hi}) is not divisible into {world} slices of whole 16-byte vectors"
        self.per = (hi - lo) // world
        self.params, self.pending, self.launched = params, 0, False
        self.state_off = 0
        self.goff, self.prev_user, self.done_event = lo, None, None   # stage-3 gradient arena (zero3.py)


def _assign(params, offsets, total, cap_elems):
    """Whole-parameter buckets in reverse order; each bucket spans [start of its first parameter in
    flat order, start of the following bucket) so ranges tile the flat buffer exactly."""
    groups, cur, cur_n = [], [], 0
    for i in reversed(range(len(params))):
        n = params[i].numel()
        if cur and cur_n + n > cap_elems:
            groups.append(cur)
            cur, cur_n = [], 0
        cur.append(i)
        cur_n += n
    if cur:
        groups.append(cur)
    out, hi = [], total
    for idxs in groups:  # groups go from the end of the flat buffer towards the start
        lo = min(offsets[i] for i in idxs)
        out.append((idxs, lo, hi))
        hi = lo
    return out


class ZeroOptimizer:
    """The object returned as ``optimizer``: exposes ``param_groups`` like a torch optimizer."""

    def __init__(self, engine, params, lr, betas, eps, weight_decay):
        self.engine = engine
        self.param_groups = [{"params": params, "lr": lr, "betas": betas, "eps": eps,
                              "weight_decay": weight_decay, "bias_correction": True}]

    def zero_grad(self, set_to_none=True):
        self.engine.zero_grad()

    def step(self):
        self.engine.step()

    def state_dict(self):
        return self.engine.optimizer_state_dict()

    def load_state_dict(self, sd):
        self.engine.load_optimizer_state_dict(sd)


def _reject_unsupported(config: dict) -> None:
    """The reference's ds_config uses train_micro_batch_size_per_gpu, optimizer, comms_logger and
    zero_optimization.{stage, reduce_bucket_size} (zero_dp_training.py:26-44).  DeepSpeed keys that
    would change the training semantics but are not implemented here fail loudly instead of being
    ignored."""
    z = config.get("zero_optimization", {}) or {}
    problems = []
    if int(config.get("gradient_accumulation_steps", 1) or 1) != 1:
        problems.append("gradient_accumulation_steps > 1")
    if float(config.get("gradient_clipping", 0.0) or 0.0) != 0.0:
        problems.append("gradient_clipping")
    if (config.get("fp16") or {}).get("enabled", False):
        problems.append("fp16 (use bf16)")
    for key in ("offload_optimizer", "offload_param"):
        if (z.get(key) or {}).get("device", "none") not in ("none", None):
            problems.append(f"zero_optimization.{key}")
    if (config.get("scheduler") or {}).get("type"):
        problems.append("scheduler (step a torch scheduler on engine.optimizer instead)")
    if problems:
        raise NotImplementedError("ds_config keys not supported by dtds_b200.parallel.zero: " + ", ".join(problems))


class ZeroEngine(nn.Module):
    def __init__(self, model: nn.Module, config: dict, model_parameters=None, process_group=None):
        super().__init__()
        self.config = config
        _reject_unsupported(config)
        zcfg = config.get("zero_optimization", {}) or {}
        self.stage = int(zcfg.get("stage", 0))
        assert self.stage in (0, 1, 2, 3), f"invalid ZeRO stage {self.stage}"
        self.reduce_bucket_size = int(float(zcfg.get("reduce_bucket_size", 5e8)))
        self.allreduce_bucket_size = int(float(zcfg.get("allreduce_bucket_size", 5e8)))
        self.overlap_comm = bool(zcfg.get("overlap_comm", self.stage >= 2))
        self.micro_batch = int(config.get("train_micro_batch_size_per_gpu", 1))
        comm.configure(config.get("comms_logger"))
        if not comm.is_initialized() and int(os.environ.get("WORLD_SIZE", "1")) > 1:
            comm.init_process_group()
        self.group = process_group
        self.world = comm.get_world_size(process_group)
        self.rank = comm.get_rank(process_group)
        if torch.cuda.is_available():
            self.device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
            torch.cuda.set_device(self.device)
        else:
            self.device = torch.device("cpu")
        want_bf16 = bool((config.get("bf16") or {}).get("enabled", False))
        self.module = model.to(self.device)
        if want_bf16:
            self.module = self.module.to(torch.bfloat16)
        params = [p for p in (model_parameters if model_parameters is not None else model.parameters())
                  if p.requires_grad]
        ocfg = config.get("optimizer", {}) or {}
        op = ocfg.get("params", {}) or {}
        self.opt_type = ocfg.get("type", "Adam")
        lr = float(op.get("lr", 1e-3))
        betas = tuple(op.get("betas", (0.9, 0.999)))
        eps = float(op.get("eps", 1e-8))
        wd = float(op.get("weight_decay", 0.0))
        # DeepSpeed maps "Adam" to FusedAdam with adam_w_mode=True; "AdamW" likewise.
        self.adamw_mode = bool(op.get("adam_w_mode", True))
        # every parameter offset (hence every bucket edge) and the total are multiples of
        # lcm(64, 8*world): buckets split evenly over the ranks in whole 16-byte vectors for any
        # world size (3, 5, 6, 7 included), none of the tail is left unreduced
        align = math.lcm(64, 8 * self.world)
        nvls = os.environ.get("DTDS_NVLS", "auto").lower()
        want_mc = self.world > 1 and (nvls in ("1", "force", "on") or (nvls == "auto" and self.world >= 4))
        self.flat = FlatBuffers(params, symmetric=True, group=process_group, align=align,
                                pad_total_to=align, multicast_grads=want_mc,
                                multicast_params=want_mc and self.stage < 3,
                                with_grads=self.stage < 3)   # stage 3 keeps gradients in a small arena
        self._want_mc = want_mc
        self.lowp = self.flat.param_dtype != torch.float32
        self._custom = self.flat.symmetric and comm.custom_kernels_active(self.flat.param_flat) \
            and self.flat.grad_dtype == torch.bfloat16
        self.flat.broadcast_params(0)  # _broadcast_model (zero_dp_training.py:46, I3)
        self.optimizer = ZeroOptimizer(self, self.flat.params, lr, betas, eps, wd)
        self.step_count = torch.zeros((), dtype=torch.int32, device=self.device)
        self._comm_stream = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None
        self._last_event = None
        self._callback_queued = False
        self._stepped_in_backward = False
        self.global_steps = 0
        self._build_state()
        for p in self.flat.params:
            p._on_grad_ready = self._param_ready
            p.register_post_accumulate_grad_hook(self._autograd_grad_hook)
        self.param3 = None
        if self.stage == 3:
            from .zero3 import Zero3ParamManager
            self.param3 = Zero3ParamManager(self)
        self._reset_iteration()

    # -------------------------------------------------------------------------- state layout
    def _build_state(self):
        f = self.flat
        dev = self.device
        if self.stage == 0:
            self.buckets = []
            n = f.total
            self.master = f.param_flat.float() if self.lowp else f.param_flat
            self.exp_avg = torch.zeros(n, dtype=torch.float32, device=dev)
            self.exp_avg_sq = torch.zeros(n, dtype=torch.float32, device=dev)
            return
        self.buckets = []
        self._bucket_of = {}
        off = 0
        for bi, (idxs, lo, hi) in enumerate(_assign(f.params, f.offsets, f.total, self.reduce_bucket_size)):
            b = _ZBucket(bi, lo, hi, self.world, [f.params[i] for i in idxs])
            b.state_off = off
            off += b.per
            self.buckets.append(b)
            for p in b.params:
                self._bucket_of[id(p)] = b
        # fp32 master weights + Adam moments for the owned 1/n slice of every bucket only
        self.master = torch.empty(off, dtype=torch.float32, device=dev)
        for b in self.buckets:
            s = b.lo + self.rank * b.per
            self.master[b.state_off:b.state_off + b.per].copy_(f.param_flat[s:s + b.per])
        self.exp_avg = torch.zeros(off, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(off, dtype=torch.float32, device=dev)

    def gradient_bytes(self) -> int:
        """Bytes of gradient storage resident on this rank (stage 3: the arena, not the model size)."""
        if self.param3 is not None:
            return self.param3.gradient_bytes()
        g = self.flat.grad_flat
        return 0 if g is None else g.numel() * g.element_size()

    def parameter_bytes(self) -> int:
        """Bytes of parameter storage resident on this rank between steps (stage 3: its shards)."""
        if self.param3 is not None:
            return self.param3.resident_parameter_bytes()
        return self.flat.param_flat.numel() * self.flat.param_flat.element_size()

    def memory_report(self) -> dict:
        """What this rank holds for the model between steps, in bytes (symmetric buffers included --
        they do not show in torch's allocator statistics because the library allocates them)."""
        free, total = torch.cuda.mem_get_info(self.device) if self.device.type == "cuda" else (0, 0)
        return {"stage": self.stage, "world": self.world, "parameters": self.parameter_bytes(),
                "gradients": self.gradient_bytes(), "optimizer": self.optimizer_state_bytes(),
                "device_used": total - free}

    def optimizer_state_bytes(self) -> int:
        return sum(t.numel() * t.element_size() for t in (self.master, self.exp_avg, self.exp_avg_sq))

    # -------------------------------------------------------------------------- hooks
    def _reset_iteration(self):
        self.flat.reset_grad_state()
        for p in self.flat.params:
            p.grad = None
        for b in self.buckets:
            b.pending = len(b.params)
            b.launched = False
        self._callback_queued = False
        self._stepped_in_backward = False

    def _autograd_grad_hook(self, p):
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
        if self.stage < 2:
            return
        if not self._callback_queued:
            self._callback_queued = True
            torch.autograd.Variable._execution_engine.queue_callback(self._end_of_backward)
        b = self._bucket_of[id(p)]
        b.pending -= 1
        if b.pending == 0 and not b.launched:
            self._launch_bucket(b, overlap=self.overlap_comm)

    def _end_of_backward(self):
        if self.param3 is not None and any(not b.launched for b in self.buckets):
            # stage 3: a bucket without gradients gets zeroed below, in an arena slot that later
            # blocks may have used in the meantime; their kernels must have drained it first
            self._join_comm()
        for b in self.buckets:
            if not b.launched:
                for p in b.params:
                    if not p._grad_written:
                        p.main_grad.zero_()
                        p._grad_written = True
                self._launch_bucket(b, overlap=self.overlap_comm)
        self._join_comm()
        self._stepped_in_backward = True
        self._callback_queued = False

    def _join_comm(self):
        if self._last_event is not None and self._comm_stream is not None:
            torch.cuda.current_stream(self.device).wait_event(self._last_event)
            self._last_event = None

    # -------------------------------------------------------------------------- bucket step
    def _hyper(self):
        g = self.optimizer.param_groups[0]
        return g["lr"], g["betas"][0], g["betas"][1], g["eps"], g["weight_decay"]

    def _launch_bucket(self, b: _ZBucket, overlap: bool):
        with nvtx_range(f"zero{self.stage}.bucket{b.index} rs+adam+ag {b.hi - b.lo} elts"):
            self._launch_bucket_impl(b, overlap)

    def _launch_bucket_impl(self, b: _ZBucket, overlap: bool):
        """Reduce-scatter + Adam + all-gather for one bucket (stages 1-3)."""
        b.launched = True
        f = self.flat
        lr, b1, b2, eps, wd = self._hyper()
        so = b.state_off
        master, m, v = (self.master[so:so + b.per], self.exp_avg[so:so + b.per],
                        self.exp_avg_sq[so:so + b.per])
        esz = torch.empty((), dtype=f.grad_dtype).element_size()
        nbytes = (b.hi - b.lo) * esz
        # stage 3: gradients live in the manager's arena (ring of bucket-sized slots), not in a
        # model-sized buffer
        gbuf = self.param3.grad_buf if self.param3 is not None else f.grad_buf
        goff = b.goff if self.param3 is not None else b.lo
        if self._custom and self.world > 1:
            cur = torch.cuda.current_stream(self.device)
            stream = self._comm_stream if overlap else cur
            from ..ops.functional import wgrad_side_stream
            side = wgrad_side_stream(self.device)
            if overlap:
                ready = torch.cuda.Event()
                ready.record(cur)
                stream.wait_event(ready)
            if side is not None and side != stream:
                stream.wait_stream(side)   # gradients of this bucket produced on the wgrad stream
            ctx = comm.symm_context(self.group)
            timing = comm.comms_logger.enabled and not torch.cuda.is_current_stream_capturing()
            with torch.cuda.stream(stream):
                s, e = torch.cuda.Event(enable_timing=timing), torch.cuda.Event(enable_timing=timing)
                s.record()
                shard = self.param3.shard_view(b) if self.param3 is not None else None
                ctx.zero_step(gbuf, None if self.param3 is not None else f.param_buf,
                              goff_elems=goff, poff_elems=b.lo, per_rank_elems=b.per, master=master, m=m,
                              v=v, step=self.step_count, lr=lr, beta1=b1, beta2=b2, eps=eps,
                              weight_decay=wd, grad_scale=1.0 / self.world, adamw_mode=self.adamw_mode,
                              param_shard=shard, push_params=self.param3 is None)
                e.record()
            if timing:
                comm.comms_logger.record_events("zero_step(rs+adam+ag)", nbytes, self.world, s, e)
            if overlap:
                self._last_event = e
            if self.param3 is not None:
                b.done_event = e   # the slot may be refilled only after this kernel (see zero3.py)
            return
        # ---- library path: reduce-scatter, flat Adam on the owned slice, all-gather
        g_bucket = self.param3.grad_view(b) if self.param3 is not None else f.grad_flat[b.lo:b.hi]
        if self.world > 1:
            g_own = torch.empty(b.per, dtype=g_bucket.dtype, device=self.device)
            comm.reduce_scatter_tensor(g_own, g_bucket, group=self.group)
        else:
            g_own = g_bucket
        if self.param3 is not None:   # stage 3: the owned slice lives in the shard buffer
            own = self.param3.shard_view(b)
        else:
            own = f.param_flat[b.lo + self.rank * b.per: b.lo + (self.rank + 1) * b.per]
        K.adam_flat(master, g_own, m, v, self.step_count, lr=lr, beta1=b1, beta2=b2, eps=eps,
                    weight_decay=wd, adamw_mode=self.adamw_mode, grad_scale=1.0 / self.world,
                    p_lp=own if self.lowp else None, inc_step=False)
        if not self.lowp:
            own.copy_(master)
        if self.param3 is None and self.world > 1:
            comm.all_gather_into_tensor(f.param_flat[b.lo:b.hi], own.clone(), group=self.group)

    # -------------------------------------------------------------------------- public API
    def forward(self, *args, **kwargs):
        args = tuple(a.to(self.device, non_blocking=True) if torch.is_tensor(a) else a for a in args)
        kwargs = {k: (v.to(self.device, non_blocking=True) if torch.is_tensor(v) else v)
                  for k, v in kwargs.items()}
        if torch.is_grad_enabled():
            self._reset_iteration()
        return self.module(*args, **kwargs)

    def backward(self, loss):
        loss.backward()
        if self.stage == 0 and self.world > 1:
            # engine-side bucketed all-reduce after backward, not overlapped (I4)
            f = self.flat
            f.absorb_autograd_grads()
            f.zero_unwritten_grads()
            n, step = f.total, max(self.allreduce_bucket_size // 64 * 64, 64)
            esz = f.grad_flat.element_size()
            for lo in range(0, n, step):
                hi = min(n, lo + step)
                if self._custom:
                    ctx = comm.symm_context(self.group)
                    with comm.comms_logger.record("all_reduce", (hi - lo) * esz, self.world, f.grad_flat):
                        ctx.all_reduce(f.grad_buf, lo * esz, (hi - lo) * esz, f.grad_dtype,
                                       scale=1.0 / self.world, algo=2)
                else:
                    comm.all_reduce(f.grad_flat[lo:hi], group=self.group, backend="nccl",
                                    scale=1.0 / self.world)
        return loss

    def step(self):
        f = self.flat
        if self.stage == 0:
            if self.world == 1:
                f.absorb_autograd_grads()
                f.zero_unwritten_grads()
            lr, b1, b2, eps, wd = self._hyper()
            K.adam_flat(self.master, f.grad_flat, self.exp_avg, self.exp_avg_sq, self.step_count, lr=lr,
                        beta1=b1, beta2=b2, eps=eps, weight_decay=wd, adamw_mode=self.adamw_mode,
                        p_lp=f.param_flat if self.lowp else None, inc_step=False)
        elif self.stage == 1 or not self._stepped_in_backward:
            # stage 1: every bucket at the step boundary.  (Also the stage 2/3 fallback when the
            # model produced its gradients outside of autograd hooks.)
            f.absorb_autograd_grads()
            f.zero_unwritten_grads()
            for b in self.buckets:
                if not b.launched:
                    self._launch_bucket(b, overlap=False)
            self._join_comm()
        self.step_count += 1
        self.global_steps += 1
        if self.param3 is not None:
            self.param3.release_all()

    def zero_grad(self):
        self._reset_iteration()

    def zero_optimization_stage(self) -> int:
        return self.stage

    def train(self, mode: bool = True):
        super().train(mode)
        return self

    def eval(self):
        return self.train(False)

    @property
    def local_rank(self):
        return int(os.environ.get("LOCAL_RANK", "0"))

    def train_micro_batch_size_per_gpu(self):
        return self.micro_batch

    # -------------------------------------------------------------------------- checkpointing
    def consolidated_fp32_state_dict(self) -> dict:
        """Full fp32 ``state_dict`` on every rank: master partitions are all-gathered (stages 1-3)."""
        f = self.flat
        full = torch.empty(f.total, dtype=torch.float32, device=self.device)
        if self.stage == 0:
            full.copy_(self.master)
        else:
            for b in self.buckets:
                own = self.master[b.state_off:b.state_off + b.per]
                if self.world > 1:
                    comm.all_gather_into_tensor(full[b.lo:b.hi], own.contiguous(), group=self.group,
                                                log_name="ckpt_all_gather")
                else:
                    full[b.lo:b.hi].copy_(own)
        out = {}
        names = {id(p): n for n, p in self.module.named_parameters()}
        for p, o in zip(f.params, f.offsets):
            out[names[id(p)]] = full[o:o + p._numel].view(p._orig_shape).clone()
        for n, buf in self.module.named_buffers():
            out[n] = buf.detach().clone()
        return out

    def optimizer_state_dict(self) -> dict:
        return {"stage": self.stage, "world": self.world, "rank": self.rank,
                "master": self.master.detach().clone(), "exp_avg": self.exp_avg.clone(),
                "exp_avg_sq": self.exp_avg_sq.clone(), "step": self.step_count.clone(),
                "param_groups": [{k: v for k, v in g.items() if k != "params"}
                                 for g in self.optimizer.param_groups]}

    def load_optimizer_state_dict(self, sd: dict) -> None:
        assert sd["stage"] == self.stage and sd["world"] == self.world, "sharded state is layout-specific"
        self.master.copy_(sd["master"])
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        self.step_count.copy_(sd["step"])
        for g, s in zip(self.optimizer.param_groups, sd["param_groups"]):
            g.update(s)
        self._publish_master()

    def _publish_master(self):
        """Re-derives the working parameters from the master partitions (after a load)."""
        f = self.flat
        if self.stage == 0:
            if self.lowp:
                f.param_flat.copy_(self.master)
            return
        for b in self.buckets:
            own = self.master[b.state_off:b.state_off + b.per].to(f.param_dtype)
            if self.param3 is not None:
                self.param3.shard_view(b).copy_(own)
                continue
            if self.world > 1:
                comm.all_gather_into_tensor(f.param_flat[b.lo:b.hi], own.contiguous(), group=self.group,
                                            log_name="ckpt_all_gather")
            else:
                f.param_flat[b.lo:b.hi].copy_(own)


def initialize(model=None, model_parameters=None, config=None, optimizer=None, args=None, **kw):
    """``deepspeed.initialize`` look-alike: returns (engine, optimizer, dataloader, lr_scheduler)."""
    assert model is not None and config is not None
    engine = ZeroEngine(model, config, model_parameters=model_parameters)
    return engine, engine.optimizer, None, None
