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
                 adam_w_mode=True, bias_correction=True, grad_dtype=None):
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

    @property
    def flats(self):
        return self._flats

    def master_weights(self):
        return [g.master for g in self._groups]

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        g0 = self.param_groups[0]
        lr, (b1, b2), eps, wd = g0["lr"], g0["betas"], g0["eps"], g0["weight_decay"]
        for grp in self._groups:
            f = grp.flat
            f.absorb_autograd_grads()
            f.zero_unwritten_grads()
            dev = f.param_flat.device
            ctx = torch.cuda.device(dev) if dev.type == "cuda" else _null()
            with ctx:
                K.adam_flat(grp.master, f.grad_flat[grp.lo:grp.hi], grp.exp_avg, grp.exp_avg_sq, grp.step,
                            lr=lr, beta1=b1, beta2=b2, eps=eps, weight_decay=wd,
                            adamw_mode=self.adam_w_mode, bias_correction=g0.get("bias_correction", True),
                            p_lp=f.param_flat[grp.lo:grp.hi] if grp.lowp else None)
        return loss

    def zero_grad(self, set_to_none: bool = True):
        for f in self._flats:
            f.reset_grad_state()
            for p in f.params:
                p.grad = None

    # state_dict in torch format: per-flat fp32 master + moments + step
    def state_dict(self):
        return {"param_groups": [{k: v for k, v in g.items() if k != "params"} for g in self.param_groups],
                "flat_state": [{"master": g.master.detach().clone(), "exp_avg": g.exp_avg.clone(),
                                "exp_avg_sq": g.exp_avg_sq.clone(), "step": g.step.clone(),
                                "lo": g.lo, "hi": g.hi} for g in self._groups]}

    def load_state_dict(self, sd):
        for g, s in zip(self.param_groups, sd["param_groups"]):
            g.update(s)
        for grp, s in zip(self._groups, sd["flat_state"]):
            grp.master.copy_(s["master"])
            grp.exp_avg.copy_(s["exp_avg"])
            grp.exp_avg_sq.copy_(s["exp_avg_sq"])
            grp.step.copy_(s["step"])
            if grp.lowp:
                grp.flat.param_flat[grp.lo:grp.hi].copy_(grp.master)


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
