"""Fused Adam / AdamW over flat buffers.

One kernel launch updates every parameter (csrc/embed_loss_adam.cu: adam_flat_kernel) instead of
a Python loop over ~200-400 tensors (``transformers.AdamW`` at data_parallel_training.py:34), a
foreach multi-tensor sweep (``torch.optim.AdamW`` at model_parallel_training.py:50) or DeepSpeed's
chunked multi_tensor_adam (zero_dp_training.py:28-33).  bf16 models keep fp32 master weights and
fp32 moments here; the kernel writes the bf16 working copy back in the same pass.  The step count
lives on the device so a captured CUDA graph replays correctly.

Parameters spread over several devices (layer-split model parallel) get one flat group per device.
"""
from __future__ import annotations

import torch

from ..ops import kernels as K
from ..parallel.flat import FlatBuffers, unique_params
from ..utils.trace import nvtx_range


class _FlatGroup:
    """Optimizer state for one FlatBuffers object (or a slice [lo, hi) of it: ZeRO partition)."""

    def __init__(self, flat: FlatBuffers, lo: int = 0, hi: int | None = None):
        self.flat = flat
        self.lo, self.hi = lo, flat.total if hi is None else hi
        n = self.hi - self.lo
        dev = flat.param_flat.device
        self.lowp = flat.param_dtype != torch.float32
        if self.lowp:
            self.master = flat.param_flat[self.lo:self.hi].float()
        else:
            self.master = flat.param_flat[self.lo:self.hi]
        self.exp_avg = torch.zeros(n, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(n, dtype=torch.float32, device=dev)
        self.step = torch.zeros((), dtype=torch.int32, device=dev)


class FusedAdam(torch.optim.Optimizer):
    """Drop-in for ``torch.optim.Adam`` / ``AdamW`` (same ``param_groups`` keys: lr, betas, eps,
    weight_decay).  ``adam_w_mode=True`` (default) is decoupled weight decay."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0,
                 adam_w_mode=True, bias_correction=True, grad_dtype=None, overlap_with_backward=False,
                 overlap_bucket_mb: float = 32.0):
        params = list(params)
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay,
                        bias_correction=bias_correction)
        super().__init__(params, defaults)
        self.adam_w_mode = adam_w_mode
        self._groups: list[_FlatGroup] = []
        all_params = unique_params(p for g in self.param_groups for p in g["params"])
        # group by existing flat owner, or build one flat buffer per device for the rest
        owners, loose = {}, {}
        for p in all_params:
            fo = getattr(p, "_flat_owner", None)
            if fo is not None:
                owners[id(fo)] = fo
            else:
                loose.setdefault(p.device, []).append(p)
        self._flats = list(owners.values())
        for dev, ps in loose.items():
            self._flats.append(FlatBuffers(ps, grad_dtype=grad_dtype, device=dev))
        for f in self._flats:
            self._groups.append(_FlatGroup(f))
        # ---- optional: update each bucket as soon as its gradients are final, under backward
        self.overlap = bool(overlap_with_backward)
        self._stepped = False      # host-side: has any update run yet (see _rebind)
        self._done: dict = {}      # id(group) -> list of (lo, hi) ranges already updated this step
        self._obuckets = []        # (group, lo, hi, params)
        self._opt_streams: dict = {}
        self._overlap_bucket_mb = overlap_bucket_mb
        if self.overlap:
            self._build_overlap_buckets(overlap_bucket_mb)

    # ------------------------------------------------------------------ overlapped update
    def _build_overlap_buckets(self, bucket_mb: float) -> None:
        """Own buckets (used when no DDP engine drives the optimizer): reverse parameter order,
        `bucket_mb` MiB of parameters each, whole parameters only."""
        for grp in self._groups:
            f = grp.flat
            cap = int(bucket_mb * (1 << 20)) // f.param_flat.element_size()
            cur, cur_n = [], 0
            groups = []
            for i in reversed(range(len(f.params))):
                n = f.params[i].numel()
                if cur and cur_n + n > cap:
                    groups.append(cur)
                    cur, cur_n = [], 0
                cur.append(i)
                cur_n += n
            if cur:
                groups.append(cur)
            hi = f.total
            for idxs in groups:
                lo = min(f.offsets[i] for i in idxs)
                ps = [f.params[i] for i in idxs]
                b = {"group": grp, "lo": lo, "hi": hi, "params": ps, "pending": len(ps)}
                self._obuckets.append(b)
                hi = lo
                for p in ps:
                    p._adam_bucket = b
                    if getattr(p, "_on_grad_ready", None) is None:   # a DDP/ZeRO engine takes precedence
                        p._on_grad_ready = self._param_ready

    def _rebind(self) -> None:
        """An engine wrapped the model after this optimizer was built (the reference scripts create
        the optimizer first, data_parallel_training.py:34-44) and moved the parameters into its own
        flat buffers: follow them, carrying master weights, moments and step count over."""
        if not any(getattr(g.flat, "_superseded_by", None) is not None for g in self._groups):
            return
        new_groups = []
        for grp in self._groups:
            old = grp.flat
            new = getattr(old, "_superseded_by", None)
            if new is None:
                new_groups.append(grp)
                continue
            moved = [p for p in old.params if getattr(p, "_flat_owner", None) is new]
            stay = [p for p in old.params if getattr(p, "_flat_owner", None) is old and p.requires_grad]
            if stay or (grp.lo, grp.hi) != (0, old.total):
                raise RuntimeError("FusedAdam: only part of the parameters moved to another flat buffer; "
                                   "build the optimizer after wrapping the model")
            ng = _FlatGroup(new)
            for p in moved:
                o_old, o_new, n = old.offsets[_index_of(old.params, p)], p._flat_offset, p._numel
                ng.exp_avg[o_new:o_new + n].copy_(grp.exp_avg[o_old:o_old + n])
                ng.exp_avg_sq[o_new:o_new + n].copy_(grp.exp_avg_sq[o_old:o_old + n])
                # Masters captured at construction are only worth carrying once this optimizer has
                # updated them.  Before the first step the engine's buffer is the truth: DDP has
                # broadcast rank 0's weights into it since (the old masters still hold this rank's
                # pre-broadcast values) and _FlatGroup(new) has just derived masters from it.
                if ng.lowp and grp.lowp and self._stepped:
                    ng.master[o_new:o_new + n].copy_(grp.master[o_old:o_old + n])
            ng.step.copy_(grp.step)
            new_groups.append(ng)
            old._superseded_by = None
        self._groups = new_groups
        self._flats = [g.flat for g in new_groups]
        self._done.clear()
        if self.overlap:
            self._obuckets = []
            self._build_overlap_buckets(self._overlap_bucket_mb)

    def _param_ready(self, p) -> None:
        b = p._adam_bucket
        b["pending"] -= 1
        if b["pending"] == 0:
            self.step_range(b["group"], b["lo"], b["hi"])
        elif b["pending"] < 0:
            raise RuntimeError("FusedAdam(overlap_with_backward=True): a second backward pass ran before "
                               "zero_grad(); gradient accumulation needs overlap_with_backward=False")

    def _group_of(self, flat) -> "_FlatGroup":
        for g in self._groups:
            if g.flat is flat:
                return g
        raise KeyError("flat buffer not managed by this optimizer")

    @torch.no_grad()
    def step_range(self, group, lo: int, hi: int, stream=None) -> None:
        """Adam on flat elements [lo, hi) now (gradients of that range are final).  Runs on
        `stream` if given (a DDP engine passes its communication stream so the update follows the
        bucket's all-reduce), else on the optimizer's side stream ordered after the current stream
        and the weight-gradient stream.  ``step()`` skips ranges done here."""
        self._rebind()
        if not isinstance(group, _FlatGroup):
            group = self._group_of(group)
        f = group.flat
        dev = f.param_flat.device
        g0 = self.param_groups[0]
        lr, (b1, b2), eps, wd = g0["lr"], g0["betas"], g0["eps"], g0["weight_decay"]

        def run():
            K.adam_flat(group.master[lo - group.lo:hi - group.lo], f.grad_flat[lo:hi],
                        group.exp_avg[lo - group.lo:hi - group.lo], group.exp_avg_sq[lo - group.lo:hi - group.lo],
                        group.step, lr=lr, beta1=b1, beta2=b2, eps=eps, weight_decay=wd,
                        adamw_mode=self.adam_w_mode, bias_correction=g0.get("bias_correction", True),
                        p_lp=f.param_flat[lo:hi] if group.lowp else None, inc_step=False)

        if dev.type != "cuda":
            run()
        elif stream is not None:
            with torch.cuda.stream(stream):
                run()
            self._opt_streams[dev] = stream
        else:
            from ..ops.functional import wgrad_side_stream
            side = self._opt_streams.get(dev)
            if side is None:
                side = self._opt_streams[dev] = torch.cuda.Stream(dev)
            cur = torch.cuda.current_stream(dev)
            side.wait_stream(cur)
            ws = wgrad_side_stream(dev)
            if ws is not None:
                side.wait_stream(ws)
            with torch.cuda.stream(side):
                run()
        self._done.setdefault(id(group), []).append((lo, hi))
        self._stepped = True

    @property
    def flats(self):
        self._rebind()
        return self._flats

    def master_weights(self):
        self._rebind()
        return [g.master for g in self._groups]

    @torch.no_grad()
    def step(self, closure=None):
        with nvtx_range("FusedAdam.step"):
            return self._step(closure)

    def _step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():   # torch.optim convention: the closure recomputes loss + grads
                loss = closure()
        self._rebind()
        g0 = self.param_groups[0]
        lr, (b1, b2), eps, wd = g0["lr"], g0["betas"], g0["eps"], g0["weight_decay"]
        for grp in self._groups:
            f = grp.flat
            f.absorb_autograd_grads()
            f.zero_unwritten_grads()
            dev = f.param_flat.device
            done = sorted(self._done.pop(id(grp), []))
            if dev.type == "cuda" and dev in self._opt_streams and done:
                torch.cuda.current_stream(dev).wait_stream(self._opt_streams[dev])
            # ranges not updated under backward (all of them unless overlap is on)
            todo, pos = [], grp.lo
            for lo, hi in done:
                if lo > pos:
                    todo.append((pos, lo))
                pos = max(pos, hi)
            if pos < grp.hi:
                todo.append((pos, grp.hi))
            ctx = torch.cuda.device(dev) if dev.type == "cuda" else _null()
            with ctx:
                for lo, hi in todo:
                    K.adam_flat(grp.master[lo - grp.lo:hi - grp.lo], f.grad_flat[lo:hi],
                                grp.exp_avg[lo - grp.lo:hi - grp.lo], grp.exp_avg_sq[lo - grp.lo:hi - grp.lo],
                                grp.step, lr=lr, beta1=b1, beta2=b2, eps=eps, weight_decay=wd,
                                adamw_mode=self.adam_w_mode, bias_correction=g0.get("bias_correction", True),
                                p_lp=f.param_flat[lo:hi] if grp.lowp else None, inc_step=False)
                K.adam_flat(grp.master[:0], f.grad_flat[:0], grp.exp_avg[:0], grp.exp_avg_sq[:0], grp.step,
                            lr=lr, inc_step=True)   # advances the device-side step count only
        return loss

    def zero_grad(self, set_to_none: bool = True):
        self._rebind()
        for f in self._flats:
            f.reset_grad_state()
            for p in f.params:
                p.grad = None
        for b in self._obuckets:
            b["pending"] = len(b["params"])

    # state_dict in torch format: per-flat fp32 master + moments + step
    def state_dict(self):
        self._rebind()
        return {"param_groups": [{k: v for k, v in g.items() if k != "params"} for g in self.param_groups],
                "flat_state": [{"master": g.master.detach().clone(), "exp_avg": g.exp_avg.clone(),
                                "exp_avg_sq": g.exp_avg_sq.clone(), "step": g.step.clone(),
                                "lo": g.lo, "hi": g.hi} for g in self._groups]}

    def load_state_dict(self, sd):
        self._rebind()
        for g, s in zip(self.param_groups, sd["param_groups"]):
            g.update(s)
        flat_state = sd["flat_state"]
        if len(flat_state) != len(self._groups):
            raise ValueError(f"FusedAdam.load_state_dict: checkpoint has {len(flat_state)} flat groups, this "
                             f"optimizer has {len(self._groups)} (different wrap order or device placement?)")
        for i, (grp, s) in enumerate(zip(self._groups, flat_state)):
            want = (grp.lo, grp.hi, grp.exp_avg.numel())
            got = (s.get("lo", grp.lo), s.get("hi", grp.hi), s["exp_avg"].numel())
            if want != got:
                raise ValueError(f"FusedAdam.load_state_dict: flat group {i} covers elements [{want[0]}, {want[1]}) "
                                 f"({want[2]} values) here but [{got[0]}, {got[1]}) ({got[2]} values) in the checkpoint; "
                                 "it was written with another parameter layout or world size")
        for grp, s in zip(self._groups, flat_state):
            grp.master.copy_(s["master"])
            grp.exp_avg.copy_(s["exp_avg"])
            grp.exp_avg_sq.copy_(s["exp_avg_sq"])
            grp.step.copy_(s["step"])
            if grp.lowp:
                grp.flat.param_flat[grp.lo:grp.hi].copy_(grp.master)
        self._stepped = True   # the masters now carry trained values (see _rebind)


def _index_of(params, p) -> int:
    for i, q in enumerate(params):
        if q is p:
            return i
    raise ValueError("parameter not in flat buffer")


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


def AdamW(params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, **kw):
    """``torch.optim.AdamW`` defaults (weight_decay 1e-2)."""
    return FusedAdam(params, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, adam_w_mode=True, **kw)


def HFAdamW(params, lr=1e-3, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.0, **kw):
    """Defaults of the removed ``transformers.AdamW`` the DDP script used (eps 1e-6, wd 0)."""
    return FusedAdam(params, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, adam_w_mode=True, **kw)


@torch.no_grad()
def clip_grad_norm_(parameters, max_norm: float, eps: float = 1e-6) -> torch.Tensor:
    """``torch.nn.utils.clip_grad_norm_`` for parameters whose gradients live in flat buffers
    (``p.main_grad``; ``p.grad`` is None there, so torch's version would silently do nothing).
    One norm and one scale per flat buffer instead of per tensor; parameters without a flat
    gradient fall back to ``p.grad``.  Returns the total L2 norm before clipping.  Call it between
    ``backward()`` and ``step()`` — not with ``overlap_with_backward`` / ``fuse_optimizer``, which
    have already applied the update by then."""
    if isinstance(parameters, torch.optim.Optimizer):
        parameters = [p for g in parameters.param_groups for p in g["params"]]
    params = unique_params(parameters)
    flats, loose = {}, []
    for p in params:
        fo = getattr(p, "_flat_owner", None)
        if fo is not None and fo.grad_flat is not None:
            flats[id(fo)] = fo
        elif p.grad is not None:
            loose.append(p.grad)
    sq = None
    for fo in flats.values():
        fo.absorb_autograd_grads()
        fo.zero_unwritten_grads()
        t = fo.grad_flat.float().pow(2).sum()
        sq = t if sq is None else sq + t.to(sq.device)
    for g in loose:
        t = g.float().pow(2).sum()
        sq = t if sq is None else sq + t.to(sq.device)
    if sq is None:
        return torch.zeros(())
    total = sq.sqrt()
    scale = (max_norm / (total + eps)).clamp(max=1.0)
    for fo in flats.values():
        fo.grad_flat.mul_(scale.to(fo.grad_flat.device).to(fo.grad_flat.dtype))
    for g in loose:
        g.mul_(scale.to(g.device).to(g.dtype))
    return total
