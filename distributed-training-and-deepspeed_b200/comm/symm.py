"""Symmetric memory: identically sized device buffers on every rank, mapped into every process.

The C++ side (csrc/comm_p2p.cu) owns allocation and CUDA-IPC export/import; this module does the
handle exchange over the bootstrap process group (NCCL/gloo object collectives) and hands out
torch tensors that alias the local buffer, so GEMMs and optimizers write straight into memory
that peer GPUs can read (no staging copies).  Replaces ProcessGroupNCCL's internal buffers
(SURVEY.md N2).
"""
from __future__ import annotations

import ctypes

import torch
import torch.distributed as dist

from ..ops import _lib

DTYPE_CODE = {torch.bfloat16: 0, torch.float32: 1, torch.int64: 2}


class _RawCudaMemory:
    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 3, "strides": None}


def _all_gather_handles(handle: bytes, group) -> list[bytes]:
    world = dist.get_world_size(group)
    out = [None] * world
    dist.all_gather_object(out, handle, group=group)
    return out


class SymmBuffer:
    """One symmetric allocation of `nbytes` per rank, with the peer address table."""

    def __init__(self, nbytes: int, group=None):
        assert torch.cuda.is_available(), "symmetric memory needs CUDA"
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.nbytes = (int(nbytes) + 255) // 256 * 256
        self.device = torch.device("cuda", torch.cuda.current_device())
        L = _lib.lib()
        p = ctypes.c_ulonglong(0)
        _lib.check(L.dtds_symm_alloc(self.nbytes, ctypes.byref(p)), "dtds_symm_alloc")
        self.local_ptr = p.value
        h = ctypes.create_string_buffer(64)
        _lib.check(L.dtds_symm_export(self.local_ptr, h), "dtds_symm_export")
        handles = _all_gather_handles(h.raw, group)
        ptrs = []
        for r, hb in enumerate(handles):
            if r == self.rank:
                ptrs.append(self.local_ptr)
                continue
            q = ctypes.c_ulonglong(0)
            _lib.check(L.dtds_symm_import(hb, ctypes.byref(q)), "dtds_symm_import")
            ptrs.append(q.value)
        self.ptrs = ptrs
        self.ptr_array = (ctypes.c_ulonglong * self.world)(*ptrs)
        self._raw = _RawCudaMemory(self.local_ptr, self.nbytes)
        self._bytes = torch.as_tensor(self._raw, device=self.device)
        dist.barrier(group=group)

    def tensor(self, dtype: torch.dtype, numel: int | None = None, offset_bytes: int = 0) -> torch.Tensor:
        """Tensor aliasing the local buffer, starting `offset_bytes` in."""
        esz = torch.empty((), dtype=dtype).element_size()
        if numel is None:
            numel = (self.nbytes - offset_bytes) // esz
        assert offset_bytes % 16 == 0 and offset_bytes + numel * esz <= self.nbytes
        return self._bytes[offset_bytes:offset_bytes + numel * esz].view(dtype)

    def offset_of(self, t: torch.Tensor) -> int:
        off = t.data_ptr() - self.local_ptr
        assert 0 <= off <= self.nbytes, "tensor does not live in this symmetric buffer"
        return off


class SymmContext:
    """Per-process-group state: the signal pad plus a scratch buffer for staged collectives."""

    _by_group: dict = {}

    def __init__(self, group=None, scratch_bytes: int = 64 << 20):
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.signal = SymmBuffer(_lib.lib().dtds_signal_pad_bytes(), group)
        self._scratch = None
        self._scratch_bytes = scratch_bytes

    @classmethod
    def get(cls, group=None) -> "SymmContext":
        key = id(group) if group is not None else None
        ctx = cls._by_group.get(key)
        if ctx is None:
            ctx = cls(group)
            cls._by_group[key] = ctx
        return ctx

    def scratch(self, nbytes: int) -> SymmBuffer:
        if self._scratch is None or self._scratch.nbytes < nbytes:
            self._scratch = SymmBuffer(max(nbytes, self._scratch_bytes), self.group)
        return self._scratch

    def new_buffer(self, nbytes: int) -> SymmBuffer:
        return SymmBuffer(nbytes, self.group)

    # ---- collectives on a symmetric buffer (all ranks call with identical arguments) ----------
    def all_reduce(self, buf: SymmBuffer, offset_bytes: int, nbytes: int, dtype: torch.dtype, *,
                   scale: float = 1.0, out: torch.Tensor | None = None, algo: int = 0, blocks: int = 0):
        rc = _lib.lib().dtds_allreduce(buf.ptr_array, self.signal.ptr_array, self.rank, self.world,
                                       _lib.ptr(out), offset_bytes, nbytes, DTYPE_CODE[dtype],
                                       float(scale), int(algo), int(blocks), _lib.stream_ptr())
        _lib.check(rc, "dtds_allreduce")

    def reduce_scatter(self, buf: SymmBuffer, offset_bytes: int, nbytes_per_rank: int, dtype, out: torch.Tensor,
                       *, scale: float = 1.0, blocks: int = 0):
        rc = _lib.lib().dtds_reduce_scatter(buf.ptr_array, self.signal.ptr_array, self.rank, self.world,
                                            out.data_ptr(), offset_bytes, nbytes_per_rank, DTYPE_CODE[dtype],
                                            float(scale), int(blocks), _lib.stream_ptr())
        _lib.check(rc, "dtds_reduce_scatter")

    def all_gather(self, buf: SymmBuffer, offset_bytes: int, nbytes_per_rank: int, src: torch.Tensor,
                   *, blocks: int = 0):
        rc = _lib.lib().dtds_all_gather(buf.ptr_array, self.signal.ptr_array, self.rank, self.world,
                                        src.data_ptr(), offset_bytes, nbytes_per_rank, int(blocks),
                                        _lib.stream_ptr())
        _lib.check(rc, "dtds_all_gather")

    def broadcast(self, buf: SymmBuffer, offset_bytes: int, nbytes: int, root: int = 0, *, blocks: int = 0):
        rc = _lib.lib().dtds_broadcast(buf.ptr_array, self.signal.ptr_array, self.rank, self.world, int(root),
                                       offset_bytes, nbytes, int(blocks), _lib.stream_ptr())
        _lib.check(rc, "dtds_broadcast")

    def barrier(self):
        rc = _lib.lib().dtds_device_barrier(self.signal.ptr_array, self.rank, self.world, _lib.stream_ptr())
        _lib.check(rc, "dtds_device_barrier")

    def zero_step(self, grads: SymmBuffer, params: SymmBuffer | None, *, goff_elems: int, poff_elems: int,
                  per_rank_elems: int, master, m, v, step, lr, beta1, beta2, eps, weight_decay,
                  grad_scale=1.0, adamw_mode=True, bias_correction=True, param_shard=None,
                  push_params=True, blocks: int = 0):
        gmc = int(getattr(grads, "mc_ptr", 0) or 0)
        pmc = int(getattr(params, "mc_ptr", 0) or 0) if params is not None else 0
        if gmc or pmc:   # in-switch reduction of the gradient slice / replication of the updated slice
            rc = _lib.lib().dtds_zero_step_nvls(
                grads.ptr_array, self.signal.ptr_array, params.ptr_array if params is not None else None,
                self.rank, self.world, int(goff_elems), int(poff_elems), int(per_rank_elems),
                master.data_ptr(), m.data_ptr(), v.data_ptr(), _lib.ptr(param_shard), step.data_ptr(),
                float(lr), float(beta1), float(beta2), float(eps), float(weight_decay), float(grad_scale),
                int(adamw_mode), int(bias_correction), int(push_params), int(blocks), gmc, pmc, _lib.stream_ptr())
            _lib.check(rc, "dtds_zero_step_nvls")
            return
        rc = _lib.lib().dtds_zero_step(
            grads.ptr_array, self.signal.ptr_array, params.ptr_array if params is not None else None,
            self.rank, self.world, int(goff_elems), int(poff_elems), int(per_rank_elems),
            master.data_ptr(), m.data_ptr(), v.data_ptr(), _lib.ptr(param_shard), step.data_ptr(),
            float(lr), float(beta1), float(beta2), float(eps), float(weight_decay), float(grad_scale),
            int(adamw_mode), int(bias_correction), int(push_params), int(blocks), _lib.stream_ptr())
        _lib.check(rc, "dtds_zero_step")

    # ---- staged collective for arbitrary tensors (pytorch_allreduce.py style call) -------------
    def all_reduce_tensor(self, t: torch.Tensor, scale: float = 1.0) -> torch.Tensor:
        """In-place sum of `t` over ranks for a tensor that does not live in symmetric memory:
        copy into the symmetric scratch, reduce with the P2P kernel, result lands back in `t`."""
        assert t.is_cuda and t.is_contiguous() and t.dtype in DTYPE_CODE
        nbytes = t.numel() * t.element_size()
        padded = (nbytes + 15) // 16 * 16
        sc = self.scratch(padded)
        stage = sc.tensor(torch.uint8, padded)
        if padded != nbytes:
            stage[nbytes:].zero_()
        stage[:nbytes].copy_(t.view(torch.uint8).reshape(-1))
        if padded == nbytes and nbytes <= (512 << 10):
            self.all_reduce(sc, 0, padded, t.dtype, scale=scale, out=t, algo=1)
        else:
            self.all_reduce(sc, 0, padded, t.dtype, scale=scale, algo=2)
            t.view(torch.uint8).reshape(-1).copy_(stage[:nbytes])
        return t


def _staged_methods():
    """reduce-scatter / all-gather / broadcast for tensors outside symmetric memory, staged through the
    context's symmetric scratch buffer (the P2P kernels need peer-mapped operands)."""

    def reduce_scatter_tensor(self, output: torch.Tensor, input: torch.Tensor):
        nb = input.numel() * input.element_size()
        per = nb // self.world
        assert per * self.world == nb and per % 16 == 0 and output.numel() * output.element_size() == per
        sc = self.scratch(nb)
        sc.tensor(torch.uint8, nb).copy_(input.view(torch.uint8).reshape(-1))
        self.reduce_scatter(sc, 0, per, input.dtype, output)
        return output

    def all_gather_tensor(self, output: torch.Tensor, input: torch.Tensor):
        per = input.numel() * input.element_size()
        nb = per * self.world
        assert per % 16 == 0 and output.numel() * output.element_size() == nb
        sc = self.scratch(nb)
        self.all_gather(sc, 0, per, input)     # pushes into slot `rank` of every peer's scratch
        # the next collective on this scratch starts with a cross-GPU entry barrier, which every rank
        # reaches (in stream order) only after this copy: no peer can refill the scratch under it
        output.view(torch.uint8).reshape(-1).copy_(sc.tensor(torch.uint8, nb))
        return output

    def broadcast_tensor(self, t: torch.Tensor, src: int):
        nb = t.numel() * t.element_size()
        padded = (nb + 15) // 16 * 16
        sc = self.scratch(padded)
        stage = sc.tensor(torch.uint8, padded)
        if self.rank == src:
            stage[:nb].copy_(t.view(torch.uint8).reshape(-1))
        self.broadcast(sc, 0, padded, root=src)
        if self.rank != src:
            t.view(torch.uint8).reshape(-1).copy_(stage[:nb])
        return t

    SymmContext.reduce_scatter_tensor = reduce_scatter_tensor
    SymmContext.all_gather_tensor = all_gather_tensor
    SymmContext.broadcast_tensor = broadcast_tensor


_staged_methods()


class McastBuffer:
    """Symmetric buffer that is also bound to an NVSwitch multicast object (NVLS).

    Allocation (cuMemCreate), the exchange of shareable handles between the rank processes (file
    descriptors over unix sockets) and the cuMulticastCreate / AddDevice / BindMem sequence are the
    library's own (csrc/symm_mcast.cu); this class only sequences the steps over the bootstrap group.
    ``DTDS_MC_RUNTIME=torch`` takes the buffer from ``torch.distributed._symmetric_memory`` instead
    (also the automatic fallback if the own path fails on a platform).  ``available`` is False when
    there is no multicast address (no NVSwitch, driver without NVLS): callers stay on the P2P kernels."""

    def __init__(self, nbytes: int, group=None):
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.nbytes = (int(nbytes) + 255) // 256 * 256
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.available = False
        self.mc_ptr = 0
        self.runtime = None
        self.error = None
        self._t = None
        self._id = None
        import os
        want = os.environ.get("DTDS_MC_RUNTIME", "own").lower()
        if want != "torch":
            self._init_own()
        if not self.available and want != "own-only":
            self._init_torch()

    # ---- own runtime ---------------------------------------------------------------------------
    def _agree(self, ok: bool) -> bool:
        """True only if the step succeeded on every rank (a collective: keeps ranks in lockstep)."""
        flags = [None] * self.world
        dist.all_gather_object(flags, bool(ok), group=self.group)
        return all(flags)

    def _init_own(self):
        L = _lib.lib()
        try:
            if not self._agree(bool(L.dtds_mc_supported())):
                self.error = "multicast not supported on every device"
                return
            out_bytes = ctypes.c_size_t(0)
            bid = L.dtds_mc_create(self.nbytes, self.rank, self.world, ctypes.byref(out_bytes))
            if not self._agree(bid >= 0):
                self.error = f"dtds_mc_create -> {bid}"
                return
            self._id = bid
            import os
            info = [None] * self.world
            dist.all_gather_object(info, (os.getpid(), bid), group=self.group)   # sockets are bound by now
            rc = L.dtds_mc_root_create(bid) if self.rank == 0 else 0
            if not self._agree(rc == 0):
                self.error = f"cuMulticastCreate -> {rc}"
                return
            pids = (ctypes.c_int * self.world)(*[i[0] for i in info])
            ids = (ctypes.c_int * self.world)(*[i[1] for i in info])
            rc = L.dtds_mc_send_fds(bid, pids, ids)
            rc = rc or L.dtds_mc_recv_add(bid)
            if not self._agree(rc == 0):      # also the barrier "every device has been added"
                self.error = f"fd exchange / cuMulticastAddDevice -> {rc}"
                return
            ptrs = (ctypes.c_ulonglong * self.world)()
            mc = ctypes.c_ulonglong(0)
            rc = L.dtds_mc_bind_map(bid, ptrs, ctypes.byref(mc))
            if not self._agree(rc == 0):      # also the barrier "every rank is bound and zero-filled"
                self.error = f"cuMulticastBindMem / map -> {rc}"
                return
            self.nbytes_alloc = int(out_bytes.value)
            self.ptrs = [int(x) for x in ptrs]
            self.local_ptr = self.ptrs[self.rank]
            self.ptr_array = (ctypes.c_ulonglong * self.world)(*self.ptrs)
            self.mc_ptr = int(mc.value)
            self._raw = _RawCudaMemory(self.local_ptr, self.nbytes)
            self._t = torch.as_tensor(self._raw, device=self.device)
            self.available = self.mc_ptr != 0
            self.runtime = "own"
        except Exception as e:  # noqa: BLE001
            self.error = f"{type(e).__name__}: {e}"

    # ---- torch runtime (fallback) --------------------------------------------------------------
    def _init_torch(self):
        try:
            import torch.distributed._symmetric_memory as tsm
            self._t = tsm.empty(self.nbytes, dtype=torch.uint8, device=self.device)
            pg = self.group if self.group is not None else dist.group.WORLD
            self._hdl = tsm.rendezvous(self._t, pg)
            self.mc_ptr = int(getattr(self._hdl, "multicast_ptr", 0) or 0)
            self.ptrs = [int(p) for p in self._hdl.buffer_ptrs]
            self.local_ptr = self.ptrs[self.rank]
            self.ptr_array = (ctypes.c_ulonglong * self.world)(*self.ptrs)
            self.available = self.mc_ptr != 0
            self.runtime = "torch"
        except Exception as e:  # noqa: BLE001
            self.error = f"{self.error}; torch: {type(e).__name__}: {e}"
            self._t = None

    def tensor(self, dtype: torch.dtype, numel: int | None = None, offset_bytes: int = 0) -> torch.Tensor:
        esz = torch.empty((), dtype=dtype).element_size()
        if numel is None:
            numel = (self.nbytes - offset_bytes) // esz
        return self._t[offset_bytes:offset_bytes + numel * esz].view(dtype)

    def offset_of(self, t: torch.Tensor) -> int:
        off = t.data_ptr() - self.local_ptr
        assert 0 <= off <= self.nbytes, "tensor does not live in this symmetric buffer"
        return off


def all_reduce_nvls(ctx: "SymmContext", buf: McastBuffer, offset_bytes: int, nbytes: int, dtype, *,
                    scale: float = 1.0, blocks: int = 0) -> None:
    """In-place NVLS all-reduce of a range of a multicast-bound buffer."""
    assert buf.available and dtype in (torch.bfloat16, torch.float32)
    rc = _lib.lib().dtds_allreduce_nvls(buf.mc_ptr, ctx.signal.ptr_array, ctx.rank, ctx.world, offset_bytes,
                                        nbytes, DTYPE_CODE[dtype], float(scale), int(blocks), _lib.stream_ptr())
    _lib.check(rc, "dtds_allreduce_nvls")


def reduce_scatter_nvls(ctx: "SymmContext", buf: McastBuffer, offset_bytes: int, nbytes_per_rank: int, dtype,
                        out: torch.Tensor, *, scale: float = 1.0, blocks: int = 0) -> None:
    assert buf.available and dtype in (torch.bfloat16, torch.float32)
    rc = _lib.lib().dtds_reduce_scatter_nvls(buf.mc_ptr, ctx.signal.ptr_array, ctx.rank, ctx.world, out.data_ptr(),
                                             offset_bytes, nbytes_per_rank, DTYPE_CODE[dtype], float(scale),
                                             int(blocks), _lib.stream_ptr())
    _lib.check(rc, "dtds_reduce_scatter_nvls")


def all_gather_nvls(ctx: "SymmContext", buf: McastBuffer, offset_bytes: int, nbytes_per_rank: int,
                    src: torch.Tensor, *, blocks: int = 0) -> None:
    assert buf.available
    rc = _lib.lib().dtds_all_gather_nvls(buf.mc_ptr, ctx.signal.ptr_array, ctx.rank, ctx.world, src.data_ptr(),
                                         offset_bytes, nbytes_per_rank, int(blocks), _lib.stream_ptr())
    _lib.check(rc, "dtds_all_gather_nvls")


def broadcast_nvls(ctx: "SymmContext", buf: McastBuffer, offset_bytes: int, nbytes: int, root: int = 0, *,
                   blocks: int = 0) -> None:
    assert buf.available
    rc = _lib.lib().dtds_broadcast_nvls(buf.mc_ptr, buf.ptr_array, ctx.signal.ptr_array, ctx.rank, ctx.world,
                                        int(root), offset_bytes, nbytes, int(blocks), _lib.stream_ptr())
    _lib.check(rc, "dtds_broadcast_nvls")
