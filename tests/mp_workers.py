"""Functions executed inside spawned ranks (see tests/_mp.py).  Each takes (rank, world, *args)."""
import copy
import math

import torch
import torch.distributed as dist


def _dev(rank):
    return torch.device("cuda", rank) if torch.cuda.is_available() and dist.get_backend() == "nccl" else torch.device("cpu")


# --------------------------------------------------------------------------- collectives
def allreduce_demo(rank, world):
    """pytorch_allreduce.py:21-33: distinct int64 triples per rank, summed everywhere."""
    from dtds_b200 import comm
    base = {0: [1, 2, 3], 1: [10, 20, 30], 2: [4, 5, 6]}
    t = torch.tensor(base.get(rank, [rank, rank, rank]), device=_dev(rank))
    comm.all_reduce(t)
    want = [sum(base.get(r, [r, r, r])[i] for r in range(world)) for i in range(3)]
    assert t.tolist() == want, (t.tolist(), want)


def collectives_match_torch(rank, world):
    """Custom P2P kernels vs torch.distributed on the same inputs, many sizes and dtypes."""
    from dtds_b200 import comm
    dev = _dev(rank)
    ctx = comm.symm_context()
    torch.manual_seed(100 + rank)
    for dtype in (torch.bfloat16, torch.float32, torch.int64):
        for n in (1, 3, 8, 1000, 4096, 65536 + 8, 1 << 20, (1 << 22) + 24):
            x = torch.randint(-50, 50, (n,), device=dev).to(dtype) if dtype == torch.int64 \
                else torch.randn(n, device=dev).to(dtype)
            want = x.clone().float() if dtype != torch.int64 else x.clone()
            dist.all_reduce(want)
            got = x.clone()
            comm.all_reduce(got)
            if dtype == torch.int64:
                assert torch.equal(got, want), (dtype, n)
            else:
                tol = 2e-2 if dtype == torch.bfloat16 else 1e-5
                err = (got.float() - want.float()).abs().max().item()
                assert err <= tol * (want.abs().max().item() + 1), (dtype, n, err)
    # symmetric-buffer primitives
    n_per = 4096 + 64
    buf = ctx.new_buffer(max(n_per * world * 4, (1 << 16) * 2))
    t = buf.tensor(torch.float32, n_per * world)
    t.copy_(torch.arange(n_per * world, device=dev, dtype=torch.float32) * (rank + 1))
    torch.cuda.synchronize(); dist.barrier()
    out = torch.empty(n_per, device=dev)
    ctx.reduce_scatter(buf, 0, n_per * 4, torch.float32, out, scale=0.5)
    tot = sum(r + 1 for r in range(world))
    want = torch.arange(rank * n_per, (rank + 1) * n_per, device=dev, dtype=torch.float32) * tot * 0.5
    assert torch.allclose(out, want), "reduce_scatter"
    src = torch.full((n_per,), float(rank + 1), device=dev)
    ctx.all_gather(buf, 0, n_per * 4, src)
    torch.cuda.synchronize()
    for r in range(world):
        assert torch.all(t[r * n_per:(r + 1) * n_per] == r + 1), "all_gather"
    t.fill_(float(rank + 7))
    torch.cuda.synchronize(); dist.barrier()
    ctx.broadcast(buf, 0, n_per * world * 4, root=world - 1)
    torch.cuda.synchronize()
    assert torch.all(t == world - 1 + 7), "broadcast"
    # stress: many back-to-back in-place all-reduces reuse the signal pads without host syncs
    g = buf.tensor(torch.bfloat16, 1 << 16)
    for it in range(10000):
        g.fill_(1.0)
        ctx.all_reduce(buf, 0, (1 << 16) * 2, torch.bfloat16, scale=1.0 / world)
    torch.cuda.synchronize()
    assert torch.all(g == 1.0), "repeated all_reduce"


def zero_step_matches_reference(rank, world):
    """Fused reduce-scatter + Adam + all-gather kernel vs all-reduce + full Adam."""
    from dtds_b200 import comm
    from dtds_b200.ops import reference as ref
    dev = _dev(rank)
    ctx = comm.symm_context()
    per = 8 * 1024
    n = per * world
    gbuf, pbuf = ctx.new_buffer(n * 2), ctx.new_buffer(n * 2)
    g, p = gbuf.tensor(torch.bfloat16, n), pbuf.tensor(torch.bfloat16, n)
    torch.manual_seed(7)
    p0 = torch.randn(n, device=dev)
    p.copy_(p0)
    master = p.float()[rank * per:(rank + 1) * per].clone()
    m, v = torch.zeros(per, device=dev), torch.zeros(per, device=dev)
    step = torch.zeros((), dtype=torch.int32, device=dev)
    ref_p, ref_m, ref_v = p.float().clone(), torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    for it in range(3):
        torch.manual_seed(1000 + 10 * it + rank)
        g.copy_(torch.randn(n, device=dev))
        gsum = g.float().clone()
        dist.all_reduce(gsum)
        gavg = gsum / world
        torch.cuda.synchronize(); dist.barrier()
        ctx.zero_step(gbuf, pbuf, goff_elems=0, poff_elems=0, per_rank_elems=per, master=master, m=m, v=v,
                      step=step, lr=1e-2, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0,
                      grad_scale=1.0 / world)
        step += 1
        ref.adam_step(ref_p, gavg, ref_m, ref_v, it + 1, lr=1e-2, beta1=0.9, beta2=0.999, eps=1e-8,
                      weight_decay=0.0)
        torch.cuda.synchronize(); dist.barrier()
        err = (p.float() - ref_p).abs().max().item()
        assert err < 2e-2, f"iter {it}: params differ by {err}"
        own = (master - ref_p[rank * per:(rank + 1) * per]).abs().max().item()
        assert own < 2e-3, f"iter {it}: master partition differs by {own}"


def nvls_collectives(rank, world, runtime="own-only"):
    """Buffers bound to a multicast object by the library's own runtime (cuMem + cuMulticast, fds over
    unix sockets): multimem all-reduce / reduce-scatter / all-gather / broadcast and the ZeRO step
    kernel with in-switch reduction, against torch.distributed / the fp32 reference."""
    import os
    os.environ["DTDS_MC_RUNTIME"] = runtime
    from dtds_b200 import comm
    from dtds_b200.comm import symm
    from dtds_b200.ops import reference as ref
    dev = _dev(rank)
    ctx = comm.symm_context()
    n_per = 8192 + 64
    n = n_per * world
    buf = symm.McastBuffer(n * 4, None)
    assert buf.available, f"multicast buffer unavailable: {buf.error}"
    assert buf.runtime == ("own" if runtime.startswith("own") else "torch"), buf.runtime
    # all-reduce (fp32 and bf16)
    t = buf.tensor(torch.float32, n)
    torch.manual_seed(50 + rank)
    x = torch.randn(n, device=dev)
    t.copy_(x)
    want = x.clone(); dist.all_reduce(want)
    torch.cuda.synchronize(); dist.barrier()
    symm.all_reduce_nvls(ctx, buf, 0, n * 4, torch.float32, scale=0.5)
    torch.cuda.synchronize()
    assert torch.allclose(t, want * 0.5, atol=1e-5, rtol=1e-5), "nvls all_reduce fp32"
    # reduce-scatter
    t.copy_(torch.arange(n, device=dev, dtype=torch.float32) * (rank + 1))
    torch.cuda.synchronize(); dist.barrier()
    out = torch.empty(n_per, device=dev)
    symm.reduce_scatter_nvls(ctx, buf, 0, n_per * 4, torch.float32, out, scale=0.5)
    tot = sum(r + 1 for r in range(world))
    want = torch.arange(rank * n_per, (rank + 1) * n_per, device=dev, dtype=torch.float32) * tot * 0.5
    torch.cuda.synchronize()
    assert torch.allclose(out, want), "nvls reduce_scatter"
    # all-gather
    src = torch.full((n_per,), float(rank + 1), device=dev)
    symm.all_gather_nvls(ctx, buf, 0, n_per * 4, src)
    torch.cuda.synchronize()
    for r in range(world):
        assert torch.all(t[r * n_per:(r + 1) * n_per] == r + 1), "nvls all_gather"
    # broadcast
    t.fill_(float(rank + 7))
    torch.cuda.synchronize(); dist.barrier()
    symm.broadcast_nvls(ctx, buf, 0, n * 4, root=world - 1)
    torch.cuda.synchronize()
    assert torch.all(t == world - 1 + 7), "nvls broadcast"
    # ZeRO step with the switch reducing the gradient slice and replicating the updated slice
    per = 8 * 1024
    m_n = per * world
    gbuf, pbuf = symm.McastBuffer(m_n * 2, None), symm.McastBuffer(m_n * 2, None)
    assert gbuf.available and pbuf.available
    g, p = gbuf.tensor(torch.bfloat16, m_n), pbuf.tensor(torch.bfloat16, m_n)
    torch.manual_seed(7)
    p.copy_(torch.randn(m_n, device=dev))
    master = p.float()[rank * per:(rank + 1) * per].clone()
    m, v = torch.zeros(per, device=dev), torch.zeros(per, device=dev)
    step = torch.zeros((), dtype=torch.int32, device=dev)
    ref_p, ref_m, ref_v = p.float().clone(), torch.zeros(m_n, device=dev), torch.zeros(m_n, device=dev)
    for it in range(3):
        torch.manual_seed(1000 + 10 * it + rank)
        g.copy_(torch.randn(m_n, device=dev))
        gsum = g.float().clone(); dist.all_reduce(gsum)
        torch.cuda.synchronize(); dist.barrier()
        ctx.zero_step(gbuf, pbuf, goff_elems=0, poff_elems=0, per_rank_elems=per, master=master, m=m, v=v,
                      step=step, lr=1e-2, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, grad_scale=1.0 / world)
        step += 1
        ref.adam_step(ref_p, gsum / world, ref_m, ref_v, it + 1, lr=1e-2, beta1=0.9, beta2=0.999, eps=1e-8,
                      weight_decay=0.0)
        torch.cuda.synchronize(); dist.barrier()
        err = (p.float() - ref_p).abs().max().item()
        assert err < 2e-2, f"iter {it}: params differ by {err}"
    # the public API on ordinary tensors goes through the custom kernels too
    from dtds_b200.ops import _lib
    c0 = dict(_lib.call_counts)
    a = torch.randn(world * 1000, device=dev)
    o = torch.empty(1000, device=dev)
    comm.reduce_scatter_tensor(o, a)
    w = a.clone(); dist.all_reduce(w)
    assert torch.allclose(o, w[rank * 1000:(rank + 1) * 1000], atol=1e-4), "reduce_scatter_tensor"
    gat = torch.empty(world * 1000, device=dev)
    comm.all_gather_into_tensor(gat, o)
    assert torch.allclose(gat, w, atol=1e-4), "all_gather_into_tensor"
    b = torch.full((777,), float(rank), device=dev, dtype=torch.bfloat16)
    comm.broadcast(b, src=world - 1)
    assert torch.all(b == world - 1), "broadcast"
    used = {k for k, v in _lib.call_counts.items() if v > c0.get(k, 0)}
    assert {"dtds_reduce_scatter", "dtds_all_gather", "dtds_broadcast"} <= used, used


# --------------------------------------------------------------------------- DDP
def _tiny_bert(dev, dtype, seed=0, dropout=0.0):
    from dtds_b200.models import build_model, get_config
    torch.manual_seed(seed)
    cfg = get_config("bert-tiny").replace(hidden_dropout_prob=dropout, attention_probs_dropout_prob=dropout)
    return cfg, build_model(cfg).to(device=dev, dtype=dtype)


def ddp_matches_single_process(rank, world, bucket_mb=0.05):
    """Averaged DDP gradients == gradients of the same model on the concatenated batch."""
    from dtds_b200.parallel.ddp import DistributedDataParallel
    dev = _dev(rank)
    dtype = torch.bfloat16 if dev.type == "cuda" else torch.float32
    cfg, model = _tiny_bert(dev, dtype, seed=rank)  # different init per rank: ctor must broadcast
    ddp = DistributedDataParallel(model, bucket_cap_mb=bucket_mb)
    assert len(ddp.buckets) > 2
    _, ref_model = _tiny_bert(dev, dtype, seed=0)
    for a, b in zip(model.parameters(), ref_model.parameters()):
        assert torch.equal(a.detach().cpu(), b.detach().cpu()), "constructor did not broadcast rank 0"
    B, S = 2, 128
    g = torch.Generator().manual_seed(5)
    ids_all = torch.randint(0, cfg.vocab_size, (world * B, S), generator=g)
    labels_all = ids_all.clone()
    labels_all[:, ::2] = -100
    ids, labels = ids_all[rank * B:(rank + 1) * B].to(dev), labels_all[rank * B:(rank + 1) * B].to(dev)
    loss = ddp(ids, labels=labels).loss
    loss.backward()
    if dev.type == "cuda":
        torch.cuda.synchronize()
    ref_loss = ref_model(ids_all.to(dev), labels=labels_all.to(dev)).loss
    ref_loss.backward()
    for (n, p), q in zip(model.named_parameters(), ref_model.parameters()):
        got, want = p.main_grad.float(), q.grad.float()
        denom = want.abs().max().item() + 1e-6
        err = (got - want).abs().max().item() / denom
        tol = 0.06 if dtype == torch.bfloat16 else 1e-4
        assert err < tol, f"{n}: {err}"
    # second iteration reuses buckets and signal pads
    ddp(ids, labels=labels).loss.backward()
    if dev.type == "cuda":
        torch.cuda.synchronize()
    w = model.embeddings.word_embeddings
    g1 = w.main_grad.float().clone()
    want = ref_model.embeddings.word_embeddings.grad.float()
    assert (g1 - want).abs().max().item() / (want.abs().max().item() + 1e-6) < (0.06 if dtype == torch.bfloat16 else 1e-4)
    # no_sync accumulates locally without communication, the next synchronised pass reduces the sum
    with ddp.no_sync():
        ddp(ids, labels=labels).loss.backward()
    ddp(ids, labels=labels).loss.backward()
    if dev.type == "cuda":
        torch.cuda.synchronize()
    # local pass + synchronised pass: (g_local + g_local)/world summed over ranks = 2 x averaged gradient
    got2 = w.main_grad.float()
    err2 = (got2 - 2 * want).abs().max().item() / (2 * want.abs().max().item() + 1e-6)
    assert err2 < (0.08 if dtype == torch.bfloat16 else 1e-4), err2


def optimizer_built_before_ddp_wrap(rank, world):
    """The reference builds the optimizer before wrapping the model (data_parallel_training.py:34-44).
    With low-precision parameters the optimizer captured fp32 masters of THIS rank's initial weights;
    the wrap then broadcasts rank 0's.  After the first step all replicas must still be identical and
    equal to a run where the optimizer was built after the wrap."""
    from dtds_b200.optim import FusedAdam
    from dtds_b200.parallel.ddp import DistributedDataParallel
    dev = _dev(rank)
    dtype = torch.bfloat16

    def run(opt_first):
        cfg, model = _tiny_bert(dev, dtype, seed=100 + rank)   # replicas start DIFFERENT
        if opt_first:
            opt = FusedAdam(model.parameters(), lr=1e-2)
            ddp = DistributedDataParallel(model, bucket_cap_mb=0.05)
        else:
            ddp = DistributedDataParallel(model, bucket_cap_mb=0.05)
            opt = FusedAdam(model.parameters(), lr=1e-2)
        g = torch.Generator().manual_seed(9)
        ids = torch.randint(0, cfg.vocab_size, (2, 128), generator=g).to(dev)
        labels = ids.clone()
        labels[:, ::2] = -100
        for _ in range(2):
            opt.zero_grad()
            ddp(ids, labels=labels).loss.backward()
            opt.step()
        if dev.type == "cuda":
            torch.cuda.synchronize()
        return torch.cat([p.detach().float().flatten() for p in model.parameters()]).cpu()

    a, b = run(True), run(False)
    # Adam's first steps move every weight by ~lr whatever the gradient's size, so one last-bit
    # difference in a near-zero gradient (CPU reductions change their split under load) flips a step
    # of 2*lr; the bug this guards against (masters taken from the rank's own pre-broadcast weights)
    # changes essentially every element on the non-zero ranks.
    frac = ((a - b).abs() > 1e-6).float().mean().item()
    assert frac < 0.01, f"optimizer-first differs from wrap-first in {frac:.1%} of the elements"
    parts = [torch.empty_like(a) for _ in range(world)] if dev.type != "cuda" else None
    import torch.distributed as dist
    mine = a.to(dev) if dev.type == "cuda" else a
    gathered = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine)
    for r in range(world):
        assert torch.equal(gathered[r].cpu(), a), f"replica {r} diverged from replica {rank}"


def ddp_training_loop(rank, world, use_graph=False, fuse_opt=False):
    """A few optimizer steps under DDP; all ranks must hold identical parameters afterwards.
    fuse_opt: Adam runs per bucket on the communication stream right behind the bucket's
    all-reduce (DistributedDataParallel.fuse_optimizer) and must give the same training curve."""
    from dtds_b200.optim import FusedAdam
    from dtds_b200.parallel.ddp import DistributedDataParallel
    dev = _dev(rank)
    dtype = torch.bfloat16 if dev.type == "cuda" else torch.float32
    g = torch.Generator().manual_seed(50 + rank)
    ids, labels = None, None

    def run(fused, graph, dropout):
        nonlocal ids, labels
        cfg, model = _tiny_bert(dev, dtype, seed=0, dropout=dropout)
        ddp = DistributedDataParallel(model, bucket_cap_mb=0.1)
        opt = FusedAdam(model.parameters(), lr=1e-3)
        if fused:
            ddp.fuse_optimizer(opt)
        if ids is None:
            ids = torch.randint(0, cfg.vocab_size, (2, 128), generator=g).to(dev)
            labels = ids.clone()

        def step(i, l):
            opt.zero_grad()
            loss = ddp(i, labels=l).loss
            loss.backward()
            opt.step()
            return loss

        if graph:
            from dtds_b200.parallel.graph import GraphedStep
            gs = GraphedStep(step, [ids, labels], warmup=2)
            losses = [gs(ids, labels).item() for _ in range(6)]
        else:
            losses = [step(ids, labels).item() for _ in range(6)]
        assert losses[-1] < losses[0], losses
        flat = ddp.flat.param_flat.float()
        other = flat.clone()
        dist.broadcast(other, src=0)
        assert torch.equal(flat, other), "replicas diverged"
        return losses, flat

    if not fuse_opt:
        run(False, use_graph, 0.1)
        return
    # no dropout: the two runs see the same masks anyway, but keep the comparison free of RNG state
    l_plain, w_plain = run(False, False, 0.0)
    l_fused, w_fused = run(True, use_graph, 0.0)
    if dev.type == "cpu":
        assert torch.equal(w_plain, w_fused)
    skip = 2 if use_graph else 0   # graph warm-up consumed two optimizer steps
    for a, b in zip(l_fused, l_plain[skip:]):
        assert abs(a - b) < 0.02 * max(1.0, abs(b)), (l_plain, l_fused)


# --------------------------------------------------------------------------- ZeRO
def _ds_config(stage, bf16):
    return {"train_micro_batch_size_per_gpu": 2,
            "optimizer": {"type": "Adam", "params": {"lr": 1e-3}},
            "comms_logger": {"enabled": True, "verbose": False, "prof_all": True, "debug": False},
            # small buckets and a 2-block gradient ring: stage 3 reuses every slot several times per step
            "zero_optimization": {"stage": stage, "reduce_bucket_size": 3e4, "grad_ring_blocks": 2},
            "bf16": {"enabled": bf16}}


def zero_stages_match_stage0(rank, world, arch="bloom-tiny"):
    """Parameters after k steps are the same for ZeRO-0/1/2/3 (same data, same init)."""
    from dtds_b200 import comm
    from dtds_b200.models import build_model, get_config
    from dtds_b200.parallel import zero
    dev = _dev(rank)
    bf16 = dev.type == "cuda"
    cfg = get_config(arch).replace(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    g = torch.Generator().manual_seed(11 + rank)
    ids = torch.randint(0, cfg.vocab_size, (2, 128), generator=g)
    results, mem = {}, {}
    for stage in (0, 1, 2, 3):
        torch.manual_seed(0)
        model = build_model(cfg)
        engine, opt, _, _ = zero.initialize(model=model, model_parameters=model.parameters(),
                                            config=_ds_config(stage, bf16))
        assert engine.zero_optimization_stage() == stage
        pg = opt.param_groups[0]
        assert pg["lr"] == 1e-3 and tuple(pg["betas"]) == (0.9, 0.999) and pg["eps"] == 1e-8
        engine.train()
        losses = []
        for _ in range(4):
            loss = engine(ids.to(dev), labels=ids.to(dev)).loss
            engine.backward(loss)
            engine.step()
            losses.append(loss.item())
        assert losses[-1] < losses[0], (stage, losses)
        sd = engine.consolidated_fp32_state_dict()
        results[stage] = torch.cat([v.flatten().float().cpu() for k, v in sorted(sd.items())])
        mem[stage] = engine.optimizer_state_bytes()
        if stage == 3 and bf16 and world > 1:
            from dtds_b200.ops import _lib
            assert _lib.call_counts.get("dtds_gemm_bf16_allgather", 0) > 0, "fused all-gather GEMM was not used"
            assert _lib.call_counts.get("dtds_zero_step", 0) > 0, "fused RS+Adam kernel was not used"
        if stage == 3:
            ring = [b for b in engine.buckets if b.prev_user is not None]
            assert ring, "the gradient ring was never reused: test does not cover slot hand-over"
            n_elems = sum(p._numel for p in engine.flat.params)
            pinned = sum(b.hi - b.lo for b in engine.buckets if b.prev_user is None and b.goff < ring[0].goff)
            assert engine.gradient_bytes() < (n_elems * 0.8 + pinned) * (2 if bf16 else 4), \
                ("stage-3 gradients are not sharded", engine.gradient_bytes(), n_elems)
            assert engine.param3.resident_parameter_bytes() * world <= \
                sum(p._numel for p in engine.flat.params) * (2 if bf16 else 4) * 1.2 + 65536 * world
    for stage in (1, 2, 3):
        err = (results[stage] - results[0]).abs().max().item()
        # fp32 on CPU: same math, other summation order (bucketed reduce-scatter vs one all-reduce,
        # thread count per rank); four Adam steps amplify last-bit differences of tiny gradients
        tol = 2e-2 if bf16 else 2e-4
        assert err < tol, f"stage {stage} differs from stage 0 by {err}"
        assert mem[stage] * world <= mem[0] * 1.05 + 4096 * world, (stage, mem)
    total = comm.comms_logger.total_latency_ms()
    assert total > 0 and comm.comms_logger.comms_dict, "comms logger recorded nothing"


# --------------------------------------------------------------------------- pipeline
def pipeline_matches_single_process(rank, world, chunks=4):
    """GPipe over `world` stages: gradients and loss equal the single-process full-batch run."""
    from dtds_b200.models import get_config
    from dtds_b200.models.bert_mp import BertModelWithMP
    from dtds_b200.ops import functional as Fn
    dev = _dev(rank)
    dtype = torch.bfloat16 if dev.type == "cuda" else torch.float32
    cfg = get_config("bert-tiny").replace(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0,
                                          num_hidden_layers=4)
    torch.manual_seed(0)
    bert = BertModelWithMP(cfg, dtype=dtype)
    assert bert.device_count == world and bert.distributed
    # a full single-process copy with identical weights, built without touching the process group
    torch.manual_seed(0)
    from dtds_b200.models.bert import BertEmbeddings, BertLayer, BertOnlyMLMHead
    cfg_u = cfg.replace(tie_word_embeddings=False)
    # every stage draws the same random stream: rebuild ALL modules in order, keep mine in sync
    torch.manual_seed(123)
    full = [BertEmbeddings(cfg_u)] + [BertLayer(cfg_u) for _ in range(cfg.num_hidden_layers)] + [BertOnlyMLMHead(cfg_u)]
    full = [m.to(device=dev, dtype=dtype) for m in full]
    for idx in bert._mine:
        bert._module(idx).load_state_dict(full[idx].state_dict())
    B, S = 8, 128
    g = torch.Generator().manual_seed(3)
    ids = torch.randint(0, cfg.vocab_size, (B, S), generator=g)
    labels = ids.clone()
    labels[:, ::3] = -100
    pipe = bert.to_pipeline(chunks)
    loss = pipe.forward_backward(ids, labels)
    # reference: all modules in this process, whole batch
    h = full[0](ids.to(dev))
    for m in full[1:-1]:
        h = m(h)
    ref_loss = full[-1](h, labels=labels.to(dev))
    ref_loss.backward()
    if rank == world - 1:
        assert abs(loss.item() - ref_loss.item()) < (3e-2 if dtype == torch.bfloat16 else 1e-4), (loss.item(), ref_loss.item())
    tol = 0.08 if dtype == torch.bfloat16 else 1e-3
    for idx in bert._mine:
        for (n, p), q in zip(bert._module(idx).named_parameters(), full[idx].parameters()):
            denom = q.grad.float().abs().max().item() + 1e-6
            err = (p.grad.float() - q.grad.float()).abs().max().item() / denom
            assert err < tol, f"module {idx} {n}: {err}"
    idle = bert.collect_idle_time()
    assert set(idle.keys()) == set(range(world))


def pipeline_graphs_match_eager(rank, world, chunks=4, checkpoint="never"):
    """GPipe training for 6 steps with the per-micro-batch CUDA graphs (captured at step 3) ends with
    the same parameters as 6 eager steps from the same initial state; losses agree step by step."""
    import os
    from dtds_b200.models import get_config
    from dtds_b200.models.bert_mp import BertModelWithMP
    from dtds_b200.optim import AdamW
    dev = _dev(rank)
    assert dev.type == "cuda"
    cfg = get_config("bert-tiny").replace(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, num_hidden_layers=4)
    B, S = 8, 128
    g = torch.Generator().manual_seed(3)
    batches = []
    for _ in range(6):
        ids = torch.randint(0, cfg.vocab_size, (B, S), generator=g)
        labels = ids.clone()
        labels[:, ::3] = -100
        batches.append((ids, labels))

    def train(graphs):
        os.environ["DTDS_PIPE_GRAPHS"] = "1" if graphs else "0"
        torch.manual_seed(0)
        bert = BertModelWithMP(cfg, dtype=torch.bfloat16)
        bert.train()
        pipe = bert.to_pipeline(chunks, checkpoint=checkpoint)
        opt = AdamW(list(bert.parameters()), lr=1e-3)
        losses = []
        for ids, labels in batches:
            opt.zero_grad()
            loss = pipe.forward_backward(ids, labels)
            opt.step()
            losses.append(float(loss.item()) if loss is not None else 0.0)
        torch.cuda.synchronize()
        st = bert._stage()
        assert (st._graphs is not None) == graphs, "graph capture did not happen" if graphs else "unexpected capture"
        return losses, torch.cat([p.detach().float().flatten() for p in bert.parameters()]).cpu()

    l_e, p_e = train(False)
    l_g, p_g = train(True)
    if rank == world - 1:
        assert l_e[-1] < l_e[0], l_e
        for a, b in zip(l_e, l_g):
            assert abs(a - b) < 3e-2, (l_e, l_g)
    err = (p_e - p_g).abs().max().item()
    assert err < 2e-2, f"parameters after 6 steps differ by {err} between eager and graph replay"


def pipeline_recompute_replays_dropout(rank, world, chunks=4):
    """GPipe with dropout > 0: gradients with activation recomputation ('except_last') equal the
    gradients without it ('never') when both runs start from the same dropout seed state -- the
    recomputed forward must regenerate the masks of the first forward."""
    from dtds_b200.models import get_config
    from dtds_b200.models.bert_mp import BertModelWithMP
    from dtds_b200.ops import functional as Fn
    dev = _dev(rank)
    dtype = torch.bfloat16 if dev.type == "cuda" else torch.float32
    cfg = get_config("bert-tiny").replace(hidden_dropout_prob=0.2, attention_probs_dropout_prob=0.2,
                                          num_hidden_layers=4)
    torch.manual_seed(0)
    bert = BertModelWithMP(cfg, dtype=dtype)
    bert.train()
    B, S = 8, 128
    g = torch.Generator().manual_seed(3)
    ids = torch.randint(0, cfg.vocab_size, (B, S), generator=g)
    labels = ids.clone()
    labels[:, ::3] = -100
    pipe = bert.to_pipeline(chunks)
    grads = {}
    for mode in ("never", "except_last"):
        for p in bert.parameters():
            p.grad = None
            if getattr(p, "main_grad", None) is not None:
                p.main_grad.zero_()
        Fn.manual_seed(77)
        bert._stage().checkpoint = mode
        pipe.forward_backward(ids, labels)
        grads[mode] = [(n, (p.grad if p.grad is not None else p.main_grad).float().clone())
                       for n, p in bert.named_parameters() if p.requires_grad and
                       (p.grad is not None or getattr(p, "main_grad", None) is not None)]
    assert grads["never"], "no gradients recorded"
    for (n, a), (_, b) in zip(grads["never"], grads["except_last"]):
        denom = a.abs().max().item() + 1e-6
        err = (a - b).abs().max().item() / denom
        assert err < (2e-2 if dtype == torch.bfloat16 else 1e-5), f"{n}: recompute changed the gradient by {err}"


def allgather_gemm_matches_gathered(rank, world):
    """The all-gather fused GEMM (weights pulled from peer shards by the kernel's fetch warps)
    equals the GEMM on a weight that was gathered beforehand; forward and dgrad forms."""
    from dtds_b200 import comm
    from dtds_b200.ops import kernels as K
    from dtds_b200.ops import reference as ref
    dev = _dev(rank)
    ctx = comm.symm_context()
    for (M, N, Kd) in [(512, 3072, 1024), (512, 1024, 4096), (2048, 768, 768)]:
        torch.manual_seed(3)
        w_full = (torch.randn(N, Kd, device=dev) * 0.05).to(torch.bfloat16)   # same on all ranks
        per = N * Kd // world
        shard_buf = ctx.new_buffer(per * 2)
        shard = shard_buf.tensor(torch.bfloat16, per)
        shard.copy_(w_full.reshape(-1)[rank * per:(rank + 1) * per])
        torch.cuda.synchronize(); dist.barrier()
        torch.manual_seed(10 + rank)
        x = torch.randn(M, Kd, device=dev).to(torch.bfloat16)
        bias = torch.randn(N, device=dev).to(torch.bfloat16)
        segs = [(shard_buf.ptrs[r], per * 2) for r in range(world)]
        # local-first order: tiles / k-blocks / chunks start at this rank's own shard
        spec = K.GatherSpec(segs, N * Kd * 2, dev, chunk_rows=32, row_bytes=Kd * 2, local_begin=rank * per * 2)
        w_local = torch.zeros(N, Kd, dtype=torch.bfloat16, device=dev)
        for rep in range(2):   # second call reuses flags with a higher epoch
            w_local.zero_()
            y = K.gemm_allgather(x, w_local, spec, bias=bias, act=ref.ACT_GELU_TANH)
            torch.cuda.synchronize()
            assert torch.equal(w_local, w_full), "kernel did not assemble the full weight"
            want = K.gemm(x, w_full, bias=bias, act=ref.ACT_GELU_TANH)
            assert torch.equal(y, want), (M, N, Kd, rep, (y.float() - want.float()).abs().max().item())
        # dgrad form: dx = dy · W with W consumed MN-major, flags along K
        dy = torch.randn(M, N, device=dev).to(torch.bfloat16)
        w_local.zero_()
        dx = K.gemm_allgather(dy, w_local, spec, b_mn_major=True)
        torch.cuda.synchronize()
        want = K.gemm(dy, w_full, b_mn_major=True)
        # the k-blocks are visited starting at this rank's own shard: same products, another fp32
        # summation order, so the bf16 results may differ in the last bit
        err = (dx.float() - want.float()).abs().max().item()
        assert err <= 2 ** -6 * want.float().abs().max().item(), ("dgrad", M, N, Kd, err)
        dist.barrier()


def zero_checkpoint_resume(rank, world, stage=2):
    """Save at step 2, rebuild the engine from scratch, load, continue: same parameters as the
    uninterrupted run; the consolidated model file has the module's parameter names."""
    import os
    import tempfile
    from dtds_b200.models import build_model, get_config
    from dtds_b200.parallel import zero
    from dtds_b200.utils.ckpt import load_checkpoint, save_checkpoint
    dev = _dev(rank)
    bf16 = dev.type == "cuda"
    cfg = get_config("opt-tiny").replace(hidden_dropout_prob=0.0)
    ids = torch.randint(0, cfg.vocab_size, (2, 128), generator=torch.Generator().manual_seed(21 + rank))

    def make():
        torch.manual_seed(0)
        model = build_model(cfg)
        eng, opt, _, _ = zero.initialize(model=model, model_parameters=model.parameters(), config=_ds_config(stage, bf16))
        return eng, opt

    def steps(eng, n):
        for _ in range(n):
            loss = eng(ids.to(dev), labels=ids.to(dev)).loss
            eng.backward(loss)
            eng.step()

    e1, _ = make()
    steps(e1, 4)
    want = e1.consolidated_fp32_state_dict()
    e2, o2 = make()
    steps(e2, 2)
    path = [tempfile.mkdtemp(prefix="dtds_ckpt_") if rank == 0 else None]
    dist.broadcast_object_list(path, src=0)
    save_checkpoint(path[0], e2, o2, step=2)
    e3, o3 = make()
    assert load_checkpoint(path[0], e3, o3) == 2
    steps(e3, 2)
    got = e3.consolidated_fp32_state_dict()
    assert set(got) == set(want)
    for k in want:
        assert torch.allclose(got[k].float().cpu(), want[k].float().cpu(), atol=1e-6 if not bf16 else 2e-2), k
    if rank == 0:
        sd = torch.load(os.path.join(path[0], "model.pt"))
        names = {n for n, _ in e3.module.named_parameters()}
        assert names <= set(sd.keys())


def pipeline_checkpoint_resume(rank, world, chunks=2):
    """One process per stage: the checkpoint's single model.pt holds every stage's tensors, and a
    run resumed from it continues like the uninterrupted one."""
    import os
    import tempfile
    from dtds_b200.models import get_config
    from dtds_b200.models.bert_mp import BertModelWithMP
    from dtds_b200.optim import AdamW
    from dtds_b200.utils.ckpt import load_checkpoint, save_checkpoint
    dev = _dev(rank)
    dtype = torch.bfloat16 if dev.type == "cuda" else torch.float32
    cfg = get_config("bert-tiny").replace(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0,
                                          num_hidden_layers=4)
    g = torch.Generator().manual_seed(7)
    ids = torch.randint(0, cfg.vocab_size, (4, 128), generator=g)
    labels = ids.clone()
    labels[:, ::2] = -100

    def make():
        torch.manual_seed(0)
        bert = BertModelWithMP(cfg, dtype=dtype)
        pipe = bert.to_pipeline(chunks)
        params = [p for p in bert.parameters()]
        opt = AdamW(params, lr=1e-3) if params else None
        return bert, pipe, opt

    def steps(pipe, opt, n):
        for _ in range(n):
            pipe.forward_backward(ids, labels)
            if opt is not None:
                opt.step()
                opt.zero_grad()

    b1, p1, o1 = make()
    steps(p1, o1, 4)
    want = {k: v.detach().float().cpu().clone() for k, v in b1.state_dict().items()}
    b2, p2, o2 = make()
    steps(p2, o2, 2)
    path = [tempfile.mkdtemp(prefix="dtds_pp_ckpt_") if rank == 0 else None]
    dist.broadcast_object_list(path, src=0)
    save_checkpoint(path[0], p2, o2, step=2)
    b3, p3, o3 = make()
    with torch.no_grad():   # make sure the load really restores something
        for p in b3.parameters():
            p.add_(1.0)
    assert load_checkpoint(path[0], p3, o3) == 2
    steps(p3, o3, 2)
    for k, v in b3.state_dict().items():
        assert torch.allclose(v.detach().float().cpu(), want[k], atol=1e-6 if dtype == torch.float32 else 3e-2), k
    if rank == 0:
        sd = torch.load(os.path.join(path[0], "model.pt"))
        L = cfg.num_hidden_layers
        assert any(k.startswith("embeddings.") for k in sd) and any(k.startswith("head.") for k in sd)
        assert all(any(k.startswith(f"encoders.{i}.") for k in sd) for i in range(L)), sorted(sd)[:8]
