"""Flat parameter / gradient storage shared by the DDP and ZeRO engines and the fused optimizer.

All parameters of a model are re-pointed at views of ONE contiguous buffer and get a
``main_grad`` view into ONE contiguous gradient buffer.  Consequences:
  * a DDP bucket or a ZeRO partition is just a byte range — nothing is ever packed or unpacked
    (DDP's copy-into-bucket / copy-back, ZeRO's flatten / unflatten: SURVEY.md K14, K15);
  * weight-gradient GEMMs write straight into the range that the collective kernel will reduce;
  * with ``symmetric=True`` the buffers are symmetric memory, so peer GPUs read gradients and
    write updated parameters in place over NVLink;
  * the optimizer is one kernel launch over the whole range.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def unique_params(params) -> list:
    seen, out = set(), []
    for p in params:
        if id(p) not in seen:
            seen.add(id(p))
            out.append(p)
    return out


def grad_uses(p) -> int:
    """How many backward producers must report before `p`'s gradient-ready callback fires."""
    return max(getattr(p, "_grad_uses", 1), 1)


class FlatBuffers:
    def _try_multicast(self, nbytes, group):
        """A multicast-bound symmetric buffer, or None unless every rank got one."""
        from ..comm.symm import McastBuffer
        mb = McastBuffer(nbytes, group)
        ok = torch.tensor([1 if mb.available else 0], device=self.device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
        return mb if int(ok.item()) == 1 else None

    def __init__(self, params, *, param_dtype: torch.dtype | None = None,
                 grad_dtype: torch.dtype | None = None, device=None, align: int = 64,
                 pad_total_to: int = 64, symmetric: bool = False, group=None,
                 with_grads: bool = True, multicast_grads: bool = False,
                 multicast_params: bool = False):
        self.params = unique_params(params)
        assert self.params, "no parameters"
        p0 = self.params[0]
        self.device = torch.device(device) if device is not None else p0.device
        self.param_dtype = param_dtype or p0.dtype
        self.grad_dtype = grad_dtype or self.param_dtype
        self.offsets, off = [], 0
        for p in self.params:
            off = (off + align - 1) // align * align
            self.offsets.append(off)
            off += p.numel()
        self.used = off
        self.total = (off + pad_total_to - 1) // pad_total_to * pad_total_to
        from .. import comm as _comm
        self.symmetric = symmetric and self.device.type == "cuda" and dist.is_initialized() \
            and dist.get_world_size(group) > 1 and not _comm.spans_nodes()   # peer mapping is per box
        self.group = group
        self.param_buf = self.grad_buf = None  # SymmBuffer objects when symmetric
        esz_p = torch.empty((), dtype=self.param_dtype).element_size()
        esz_g = torch.empty((), dtype=self.grad_dtype).element_size()
        if self.symmetric:
            from ..comm.symm import SymmContext
            ctx = SymmContext.get(group)
            self.param_buf = None
            if multicast_params:   # NVLS: updated parameter slices are replicated by the switch
                self.param_buf = self._try_multicast(self.total * esz_p, group)
            if self.param_buf is None:
                self.param_buf = ctx.new_buffer(self.total * esz_p)
            self.param_flat = self.param_buf.tensor(self.param_dtype, self.total)
            if with_grads:
                self.grad_buf = None
                if multicast_grads:  # NVLS: gradient buffer bound to an NVSwitch multicast object
                    self.grad_buf = self._try_multicast(self.total * esz_g, group)
                if self.grad_buf is None:
                    self.grad_buf = ctx.new_buffer(self.total * esz_g)
                self.grad_flat = self.grad_buf.tensor(self.grad_dtype, self.total)
                self.grad_flat.zero_()
        else:
            self.param_flat = torch.zeros(self.total, dtype=self.param_dtype, device=self.device)
            if with_grads:
                self.grad_flat = torch.zeros(self.total, dtype=self.grad_dtype, device=self.device)
        if not with_grads:
            self.grad_flat = None
        with torch.no_grad():
            for p, o in zip(self.params, self.offsets):
                view = self.param_flat[o:o + p.numel()].view(p.shape)
                view.copy_(p.data)
                p.data = view
                if with_grads:
                    p.main_grad = self.grad_flat[o:o + p.numel()].view(p.shape)
                prev = getattr(p, "_flat_owner", None)
                if prev is not None and prev is not self:
                    prev._superseded_by = self   # optimizers built on `prev` rebind (optim/adam.py)
                p._flat_owner = self
                p._orig_shape = tuple(p.shape)
                p._numel = p.numel()
                p._flat_offset = o
                p._grad_written = False
                p._uses_pending = grad_uses(p)

    # ------------------------------------------------------------------------------------
    def range_of(self, p) -> tuple[int, int]:
        return p._flat_offset, p._flat_offset + p.numel()

    def reset_grad_state(self) -> None:
        """Start of a new backward pass: the first gradient write of each parameter overwrites."""
        for p in self.params:
            p._grad_written = False
            p._uses_pending = grad_uses(p)

    def zero_unwritten_grads(self) -> None:
        """Parameters that received no gradient this step must not see last step's values."""
        for p in self.params:
            if not p._grad_written and p.requires_grad:
                p.main_grad.zero_()

    def absorb_autograd_grads(self) -> None:
        """Folds ``p.grad`` tensors produced by ordinary autograd ops into ``main_grad``."""
        for p in self.params:
            if p.grad is not None and p.grad.data_ptr() != p.main_grad.data_ptr():
                if p._grad_written:
                    p.main_grad.add_(p.grad.to(self.grad_dtype))
                else:
                    p.main_grad.copy_(p.grad)
                p._grad_written = True
                p.grad = None

    def broadcast_params(self, src: int = 0) -> None:
        """Rank `src`'s parameters to everyone (DDP constructor / DeepSpeed _broadcast_model:
        data_parallel_training.py:44, zero_dp_training.py:46)."""
        if not dist.is_initialized() or dist.get_world_size(self.group) == 1:
            return
        from .. import comm
        if self.symmetric and comm.custom_kernels_active(self.param_flat):
            ctx = comm.symm_context(self.group)
            nbytes = self.total * self.param_flat.element_size()
            with comm.comms_logger.record("broadcast", nbytes, ctx.world, self.param_flat):
                ctx.broadcast(self.param_buf, 0, nbytes, root=src)
        else:
            comm.broadcast(self.param_flat, src=src, group=self.group)

    def state_dict_views(self) -> dict:
        return {"param_flat": self.param_flat, "grad_flat": self.grad_flat}
