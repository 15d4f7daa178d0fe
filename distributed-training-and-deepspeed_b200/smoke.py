"""One tiny forward + backward + optimizer step of the flagship model (BERT MLM) on cuda:0,
through the native sm_100a kernels.  Called by ``__graft_entry__.smoke()``."""
from __future__ import annotations

import torch


def run_smoke(verbose: bool = True) -> float:
    from .models import build_model, get_config
    from .ops import _lib
    from .optim import FusedAdam

    assert torch.cuda.is_available()
    _lib.lib()  # fail loudly if the native library is missing
    torch.cuda.set_device(0)
    torch.manual_seed(0)
    cfg = get_config("bert-tiny")
    model = build_model(cfg).to(device="cuda", dtype=torch.bfloat16)
    opt = FusedAdam(model.parameters(), lr=1e-3)
    ids = torch.randint(0, cfg.vocab_size, (2, 128), device="cuda")
    labels = ids.clone()
    labels[:, ::2] = -100
    _lib.reset_launch_count()
    losses = []
    for _ in range(3):
        opt.zero_grad()
        loss = model(ids, labels=labels).loss
        loss.backward()
        opt.step()
        losses.append(float(loss.item()))
    torch.cuda.synchronize()
    assert all(l == l for l in losses), f"NaN loss: {losses}"
    assert losses[-1] < losses[0], f"loss did not decrease: {losses}"
    if verbose:
        print(f"smoke ok: losses {['%.4f' % l for l in losses]}, native launches {_lib.launch_count()}")
    return losses[-1]


def _multi_worker(rank: int, world: int, port: int) -> None:
    import os

    import torch.distributed as dist

    os.environ.update({"MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "RANK": str(rank),
                       "WORLD_SIZE": str(world), "LOCAL_RANK": str(rank)})
    torch.cuda.set_device(rank)
    from . import comm
    from .ops import reference as ref
    comm.init_process_group(backend="nccl", rank=rank, world_size=world)
    dev = torch.device("cuda", rank)
    ctx = comm.symm_context()
    # in-kernel all-reduce (two-shot P2P over peer memory) against NCCL
    n = 1 << 20
    buf = ctx.new_buffer(n * 2)
    t = buf.tensor(torch.bfloat16, n)
    torch.manual_seed(rank)
    x = torch.randn(n, device=dev).to(torch.bfloat16)
    t.copy_(x)
    want = x.float()
    dist.all_reduce(want)
    torch.cuda.synchronize()
    dist.barrier()
    ctx.all_reduce(buf, 0, n * 2, torch.bfloat16, scale=1.0 / world, algo=2)
    torch.cuda.synchronize()
    err = (t.float() - want / world).abs().max().item()
    assert err < 2e-2, f"custom all-reduce differs from NCCL by {err}"
    # fused reduce-scatter + Adam + all-gather kernel against all-reduce + reference Adam
    per = 8192
    m_n = per * world
    gbuf, pbuf = ctx.new_buffer(m_n * 2), ctx.new_buffer(m_n * 2)
    g, p = gbuf.tensor(torch.bfloat16, m_n), pbuf.tensor(torch.bfloat16, m_n)
    torch.manual_seed(7)
    p.copy_(torch.randn(m_n, device=dev))
    master = p.float()[rank * per:(rank + 1) * per].clone()
    m, v = torch.zeros(per, device=dev), torch.zeros(per, device=dev)
    step = torch.zeros((), dtype=torch.int32, device=dev)
    ref_p, ref_m, ref_v = p.float().clone(), torch.zeros(m_n, device=dev), torch.zeros(m_n, device=dev)
    torch.manual_seed(100 + rank)
    g.copy_(torch.randn(m_n, device=dev))
    gsum = g.float().clone()
    dist.all_reduce(gsum)
    torch.cuda.synchronize()
    dist.barrier()
    ctx.zero_step(gbuf, pbuf, goff_elems=0, poff_elems=0, per_rank_elems=per, master=master, m=m, v=v, step=step,
                  lr=1e-2, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, grad_scale=1.0 / world)
    ref.adam_step(ref_p, gsum / world, ref_m, ref_v, 1, lr=1e-2, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0)
    torch.cuda.synchronize()
    dist.barrier()
    err = (p.float() - ref_p).abs().max().item()
    assert err < 2e-2, f"fused ZeRO step differs from the reference by {err}"
    if rank == 0:
        print(f"smoke (2 ranks) ok: in-kernel all-reduce and fused reduce-scatter + Adam + all-gather match NCCL / reference")
    dist.barrier()
    dist.destroy_process_group()


def run_smoke_multi(world: int = 2) -> None:
    """2-rank checks on real GPUs: the P2P all-reduce kernel against NCCL and the one-launch ZeRO step
    against all-reduce + reference Adam.  Called by ``__graft_entry__.smoke()`` when >= 2 GPUs are visible."""
    import socket

    import torch.multiprocessing as mp

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_multi_worker, args=(world, port), nprocs=world, join=True)
