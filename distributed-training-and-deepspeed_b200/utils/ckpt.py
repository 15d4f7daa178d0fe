"""Checkpoint save / load / resume in the torch ``state_dict`` format.

The reference has no training checkpoints (only ``from_pretrained`` loads and a ``state_dict`` ->
``load_state_dict`` copy of one OPT layer, model/transformer.py:115-123).  Here:

  * the model file ``model.pt`` is a plain ``state_dict`` with this repository's parameter names
    (full fp32 tensors, consolidated from the master partitions for ZeRO 1-3), written by rank 0;
  * every rank writes ``optim_rank{r}.pt`` with its optimizer state (whole state for DDP /
    single-process, the owned partitions for ZeRO) and the step counter;
  * layer-split / pipeline models (one process per stage) gather their stages' tensors into the
    same single ``model.pt``; on load every stage picks the keys of the modules it owns;
  * ``load_checkpoint`` restores both, so a run resumed at step k continues bit-for-bit like
    the uninterrupted one (tests/test_tools_cpu.py, tests/test_dist_cpu.py).
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def _rank_world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def _unwrap(model):
    return getattr(model, "module", model)


def _stage_partitioned(model) -> bool:
    """Layer-split / pipeline models driven by one process per stage hold only their own modules
    (models/bert_mp.py: the others are placeholders), so a full state dict has to be gathered."""
    m = _unwrap(model)
    m = getattr(m, "bert", m)   # the pipeline adapter wraps the model
    return bool(getattr(m, "distributed", False)) and hasattr(m, "group_ranges")


def model_state_dict(model) -> dict:
    """Full state dict on every rank (fp32 master values when an engine / fused optimizer has them;
    gathered across stages for stage-partitioned models — a collective: every rank must call)."""
    if hasattr(model, "consolidated_fp32_state_dict"):
        return model.consolidated_fp32_state_dict()
    target = _unwrap(model)
    target = getattr(target, "bert", target) if _stage_partitioned(model) else target
    local = {k: v.detach().float().cpu() if v.is_floating_point() else v.detach().cpu()
             for k, v in target.state_dict().items()}
    if _stage_partitioned(model):
        parts = [None] * dist.get_world_size(target.group)
        dist.all_gather_object(parts, local, group=target.group)
        merged = {}
        for part in parts:
            overlap = set(part) & set(merged)
            if overlap:
                raise RuntimeError(f"stages report the same parameters: {sorted(overlap)[:4]}")
            merged.update(part)
        return merged
    return local


def save_checkpoint(path: str, model, optimizer=None, step: int = 0, extra: dict | None = None) -> None:
    rank, world = _rank_world()
    os.makedirs(path, exist_ok=True)
    sd = model_state_dict(model)  # collective for ZeRO engines: every rank must call
    if rank == 0:
        torch.save({k: v.cpu() for k, v in sd.items()}, os.path.join(path, "model.pt"))
        torch.save({"step": step, "world_size": world, "extra": extra or {}}, os.path.join(path, "meta.pt"))
    if optimizer is not None:
        osd = optimizer.state_dict()
        torch.save(_to_cpu(osd), os.path.join(path, f"optim_rank{rank}.pt"))
    if world > 1:
        dist.barrier()


def _to_cpu(obj):
    if torch.is_tensor(obj):
        return obj.detach().cpu()
    if isinstance(obj, dict):
        return {k: _to_cpu(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_to_cpu(v) for v in obj)
    return obj


def load_checkpoint(path: str, model, optimizer=None, strict: bool = True) -> int:
    """Loads model (and optimizer) state; returns the saved step."""
    rank, world = _rank_world()
    meta = torch.load(os.path.join(path, "meta.pt"), map_location="cpu", weights_only=False)
    sd = torch.load(os.path.join(path, "model.pt"), map_location="cpu", weights_only=False)
    target = _unwrap(model)
    if _stage_partitioned(model):
        target = getattr(target, "bert", target)   # each stage loads the keys of its own modules
    sharded_params = getattr(model, "param3", None) is not None
    have_optim = optimizer is not None and os.path.exists(os.path.join(path, f"optim_rank{rank}.pt"))
    if sharded_params and not have_optim:
        raise RuntimeError("loading a ZeRO-3 engine needs its optimizer shards (sharded parameters)")
    # 1. model.pt -> module, always: parameters an optimizer does not own (frozen ones, plain
    #    torch.optim optimizers that keep no master copy) and buffers come from nowhere else.
    if not sharded_params:
        with torch.no_grad():
            own = target.state_dict()
            missing = [k for k in own if k not in sd]
            if (strict or _stage_partitioned(model)) and missing:
                raise KeyError(f"checkpoint lacks {missing[:5]}")
            for k, v in own.items():
                if k in sd:
                    v.copy_(sd[k].to(v.dtype))
    # 2. optimizer shard: engine optimizers then overwrite what they own with their fp32 masters
    if have_optim:
        osd = torch.load(os.path.join(path, f"optim_rank{rank}.pt"), map_location="cpu", weights_only=False)
        optimizer.load_state_dict(osd)
    if world > 1:
        dist.barrier()
    return int(meta["step"])
