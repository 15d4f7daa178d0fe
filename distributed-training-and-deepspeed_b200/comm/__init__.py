"""Communication front end: process-group bootstrap, logged collectives, comms logger.

``torch.distributed`` (NCCL on GPUs, gloo on CPU) is the plumbing: rendezvous, object exchange,
pipeline p2p, and the CPU test path.  On GPUs the bandwidth-critical collectives (all-reduce,
reduce-scatter, all-gather, broadcast) run as hand-written P2P kernels over symmetric memory
(``symm.py`` / csrc/comm_p2p.cu); ``backend="nccl"`` can be forced per call to measure the
library baseline.  Every collective issued through this module is recorded by the comms logger —
the ``deepspeed.comm`` + ``comms_logger`` surface zero_dp_training.py:3,34-39,74,101-109 uses.
"""
from __future__ import annotations

import os
import time
from datetime import timedelta

import torch
import torch.distributed as dist

from .logger import CommsLogger

comms_logger = CommsLogger()
_use_custom = True


def configure(comms_config: dict | None = None, use_custom_kernels: bool | None = None) -> None:
    """Applies a ds_config-style ``comms_logger`` block and selects custom vs NCCL collectives."""
    global _use_custom
    if comms_config is not None:
        comms_logger.configure(comms_config)
    if use_custom_kernels is not None:
        _use_custom = use_custom_kernels


def is_initialized() -> bool:
    return dist.is_available() and dist.is_initialized()


def init_process_group(backend: str | None = None, rank: int | None = None, world_size: int | None = None,
                       master_addr: str | None = None, master_port: int | str | None = None,
                       timeout_s: int = 1800) -> None:
    """env:// rendezvous (MASTER_ADDR/MASTER_PORT/RANK/WORLD_SIZE), NCCL on CUDA else gloo.
    data_parallel_training.py:15-23 / pytorch_allreduce.py:7-15 hard-code 'nccl' and port 12355;
    here both are parameters and the address defaults to 127.0.0.1."""
    if is_initialized():
        return
    if master_addr is not None:
        os.environ["MASTER_ADDR"] = str(master_addr)
    if master_port is not None:
        os.environ["MASTER_PORT"] = str(master_port)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    rank = int(os.environ.get("RANK", 0)) if rank is None else rank
    world_size = int(os.environ.get("WORLD_SIZE", 1)) if world_size is None else world_size
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    kwargs = {}
    if backend == "nccl":
        local = int(os.environ.get("LOCAL_RANK", rank % max(torch.cuda.device_count(), 1)))
        torch.cuda.set_device(local)
        kwargs["device_id"] = torch.device("cuda", local)
    dist.init_process_group(backend=backend, rank=rank, world_size=world_size,
                            timeout=timedelta(seconds=timeout_s), **kwargs)


def destroy_process_group() -> None:
    if is_initialized():
        dist.destroy_process_group()


def get_rank(group=None) -> int:
    return dist.get_rank(group) if is_initialized() else 0


def get_world_size(group=None) -> int:
    return dist.get_world_size(group) if is_initialized() else 1


def get_local_rank() -> int:
    return int(os.environ.get("LOCAL_RANK", 0))


def custom_kernels_active(t: torch.Tensor | None = None) -> bool:
    """True when collectives should run as P2P kernels: CUDA tensors, >1 rank, not disabled."""
    if not _use_custom or not is_initialized() or get_world_size() < 2:
        return False
    if os.environ.get("DTDS_COMM", "custom").lower() == "nccl":
        return False
    if spans_nodes():
        return False   # peer-mapped memory / NVLS exist inside one NVSwitch domain only
    return torch.cuda.is_available() and (t is None or t.is_cuda)


def spans_nodes() -> bool:
    """True when the job has ranks on more than one box (launcher contract: LOCAL_WORLD_SIZE is the
    number of ranks on this box; absent, e.g. under mp.spawn, the job is local)."""
    local = os.environ.get("LOCAL_WORLD_SIZE")
    return bool(local) and is_initialized() and get_world_size() > int(local)


def symm_context(group=None):
    from .symm import SymmContext
    return SymmContext.get(group)


# ------------------------------------------------------------------------------- collectives
def _nbytes(t: torch.Tensor) -> int:
    return t.numel() * t.element_size()


def all_reduce(tensor: torch.Tensor, op=dist.ReduceOp.SUM, group=None, backend: str | None = None,
               scale: float = 1.0, log_name: str = "all_reduce"):
    """In-place sum over ranks (optionally scaled).  `backend='nccl'` forces the library path."""
    if not is_initialized() or get_world_size(group) == 1:
        if scale != 1.0:
            tensor.mul_(scale)
        return tensor
    with comms_logger.record(log_name, _nbytes(tensor), get_world_size(group), tensor):
        if backend != "nccl" and op == dist.ReduceOp.SUM and custom_kernels_active(tensor) \
                and tensor.dtype in (torch.bfloat16, torch.float32, torch.int64) and tensor.is_contiguous():
            symm_context(group).all_reduce_tensor(tensor, scale)
        else:
            dist.all_reduce(tensor, op=op, group=group)
            if scale != 1.0:
                tensor.mul_(scale)
    return tensor


def _staged_ok(t: torch.Tensor) -> bool:
    """The tensor can go through the symmetric scratch buffer and the P2P kernels."""
    return custom_kernels_active(t) and t.is_contiguous() and not torch.cuda.is_current_stream_capturing()


def broadcast(tensor: torch.Tensor, src: int = 0, group=None, log_name: str = "broadcast"):
    if not is_initialized() or get_world_size(group) == 1:
        return tensor
    with comms_logger.record(log_name, _nbytes(tensor), get_world_size(group), tensor):
        if _staged_ok(tensor):
            symm_context(group).broadcast_tensor(tensor, src)
        else:
            dist.broadcast(tensor, src=src, group=group)
    return tensor


def reduce_scatter_tensor(output: torch.Tensor, input: torch.Tensor, group=None,
                          log_name: str = "reduce_scatter_tensor"):
    if not is_initialized() or get_world_size(group) == 1:
        output.copy_(input.view_as(output))
        return output
    with comms_logger.record(log_name, _nbytes(input), get_world_size(group), input):
        if _staged_ok(input) and _staged_ok(output) and input.dtype in (torch.bfloat16, torch.float32, torch.int64) \
                and _nbytes(output) % 16 == 0:
            symm_context(group).reduce_scatter_tensor(output, input)
        elif input.is_cuda:
            dist.reduce_scatter_tensor(output, input, group=group)
        else:  # gloo has no reduce_scatter: all-reduce a copy and slice
            tmp = input.clone()
            dist.all_reduce(tmp, group=group)
            n = output.numel()
            output.copy_(tmp.view(-1)[get_rank(group) * n:(get_rank(group) + 1) * n].view_as(output))
    return output


def all_gather_into_tensor(output: torch.Tensor, input: torch.Tensor, group=None,
                           log_name: str = "all_gather_into_tensor"):
    if not is_initialized() or get_world_size(group) == 1:
        output.view(-1)[:input.numel()].copy_(input.view(-1))
        return output
    with comms_logger.record(log_name, _nbytes(output), get_world_size(group), input):
        if _staged_ok(input) and _staged_ok(output) and _nbytes(input) % 16 == 0:
            symm_context(group).all_gather_tensor(output, input)
        elif input.is_cuda:
            dist.all_gather_into_tensor(output, input, group=group)
        else:
            parts = list(output.view(get_world_size(group), -1).unbind(0))
            dist.all_gather(parts, input.reshape(-1), group=group)
    return output


def reduce(tensor: torch.Tensor, dst: int, group=None, log_name: str = "reduce"):
    if not is_initialized() or get_world_size(group) == 1:
        return tensor
    with comms_logger.record(log_name, _nbytes(tensor), get_world_size(group), tensor):
        dist.reduce(tensor, dst=dst, group=group)
    return tensor


def send(tensor: torch.Tensor, dst: int, group=None):
    with comms_logger.record("send", _nbytes(tensor), 2, tensor):
        dist.send(tensor, dst=dst, group=group)


def recv(tensor: torch.Tensor, src: int, group=None):
    with comms_logger.record("recv", _nbytes(tensor), 2, tensor):
        dist.recv(tensor, src=src, group=group)
    return tensor


def barrier(group=None, log_name: str = "barrier"):
    if not is_initialized() or get_world_size(group) == 1:
        return
    t0 = time.perf_counter()
    dist.barrier(group=group)
    comms_logger.append_host(log_name, 0, (time.perf_counter() - t0) * 1e3, get_world_size(group))


def log_summary(show_straggler: bool = False) -> None:
    """Prints the per-op table (``dist.log_summary()`` at zero_dp_training.py:109)."""
    barrier(log_name="log_summary_barrier")
    comms_logger.log_all(print_log=get_rank() == 0)
