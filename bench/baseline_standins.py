"""Declared stand-ins for the reference's ZeRO and model-parallel scripts on this image.

The reference's zero_dp_training.py needs DeepSpeed (no wheel here) and its model/bert_mp.py imports
torch.distributed.pipeline.sync (removed from torch) and relies on BertLayer returning a tuple
(transformers 4.29), so neither runs unmodified.  These are the closest library-only equivalents,
none of this repository's kernels or engines on the path:

  zero   torch FSDP over HF BloomForCausalLM (bloom-560m config, random init), bf16 mixed precision,
         AdamW lr 1.5e-4, batch 1 x 512 per rank:  --stage 0 -> NO_SHARD (DDP), 1 -> DDP +
         ZeroRedundancyOptimizer, 2 -> SHARD_GRAD_OP, 3 -> FULL_SHARD.   (torchrun, one rank per GPU)
  mp     HF BertEmbeddings + BertLayer x L + BertOnlyMLMHead (untied) placed on N devices with
         np.array_split in ONE process, activations moved with .to(device) -- the reference's naive
         layer-split forward (model/bert_mp.py:39-47,91-99) -- bf16 autocast, batch 16 x 512.
  pipe   the same modules, one process per stage, torch.distributed.pipelining ScheduleGPipe with 4
         micro-batches (the successor of the removed Pipe).                     (torchrun)

    python -m torch.distributed.run --nproc-per-node 8 bench/baseline_standins.py zero --stage 3
    python bench/baseline_standins.py mp --devices 4 [--model large]
    python -m torch.distributed.run --nproc-per-node 4 bench/baseline_standins.py pipe

Prints one JSON line (rank 0): ms/step over --steps steps after --warmup, CUDA events, max over ranks.
"""
import argparse
import json
import os
import time

import numpy as np
import torch
import torch.distributed as dist
import torch.nn as nn


def bert_cfg(model):
    from transformers import BertConfig
    big = model == "large"
    return BertConfig(vocab_size=28996, hidden_size=1024 if big else 768, num_hidden_layers=24 if big else 12,
                      num_attention_heads=16 if big else 12, intermediate_size=4096 if big else 3072)


def timed_loop(step, steps, warmup, world):
    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / steps], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


def run_zero(a):
    from torch.distributed.fsdp import FullyShardedDataParallel as FSDP, MixedPrecision, ShardingStrategy
    from transformers import BloomConfig, BloomForCausalLM
    dist.init_process_group("nccl")
    rank, world = dist.get_rank(), dist.get_world_size()
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    cfg = BloomConfig(vocab_size=250880, hidden_size=1024, n_layer=24, n_head=16)
    torch.manual_seed(0)
    model = BloomForCausalLM(cfg)
    mp = MixedPrecision(param_dtype=torch.bfloat16, reduce_dtype=torch.bfloat16, buffer_dtype=torch.bfloat16)
    if a.stage == 1:
        from torch.distributed.optim import ZeroRedundancyOptimizer
        net = nn.parallel.DistributedDataParallel(model.cuda(), device_ids=[torch.cuda.current_device()])
        opt = ZeroRedundancyOptimizer(net.parameters(), optimizer_class=torch.optim.AdamW, lr=1.5e-4)
        name = "DDP + ZeroRedundancyOptimizer (bf16 autocast)"
    else:
        strat = {0: ShardingStrategy.NO_SHARD, 2: ShardingStrategy.SHARD_GRAD_OP, 3: ShardingStrategy.FULL_SHARD}[a.stage]
        from torch.distributed.fsdp.wrap import ModuleWrapPolicy
        from transformers.models.bloom.modeling_bloom import BloomBlock
        net = FSDP(model, sharding_strategy=strat, mixed_precision=mp, device_id=torch.cuda.current_device(),
                   auto_wrap_policy=ModuleWrapPolicy({BloomBlock}), use_orig_params=True)
        opt = torch.optim.AdamW(net.parameters(), lr=1.5e-4)
        name = f"FSDP {strat.name} (bf16 mixed precision)"
    ids = torch.randint(0, cfg.vocab_size, (1, 512), device="cuda")

    def step():
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=a.stage == 1):
            loss = net(input_ids=ids, labels=ids).loss
        loss.backward()
        opt.step()
        opt.zero_grad(set_to_none=True)

    ms = timed_loop(step, a.steps, a.warmup, world)
    if rank == 0:
        print(json.dumps({"standin": "zero", "stage": a.stage, "impl": name, "model": "bloom-560m shape", "n_gpus": world,
                          "ms_per_step": ms, "samples_per_s": world / ms * 1e3,
                          "max_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30}))
    dist.destroy_process_group()


def bert_modules(model):
    from transformers.models.bert.modeling_bert import BertEmbeddings, BertLayer, BertOnlyMLMHead
    cfg = bert_cfg(model)
    cfg._attn_implementation = "sdpa"
    torch.manual_seed(0)
    return cfg, [BertEmbeddings(cfg)] + [BertLayer(cfg) for _ in range(cfg.num_hidden_layers)] + [BertOnlyMLMHead(cfg)]


class _Layer(nn.Module):
    """BertLayer returns a tensor or a tuple depending on the transformers version."""
    def __init__(self, m):
        super().__init__()
        self.m = m

    def forward(self, x):
        y = self.m(x)
        return y[0] if isinstance(y, tuple) else y


def run_mp(a):
    n = a.devices or torch.cuda.device_count()
    cfg, mods = bert_modules(a.model)
    mods = [mods[0]] + [_Layer(m) for m in mods[1:-1]] + [mods[-1]]
    groups = np.array_split(np.arange(len(mods)), n)
    devs = []
    for gi, g in enumerate(groups):
        for i in g:
            mods[i].to(f"cuda:{gi}")
            devs.append(f"cuda:{gi}")
    params = [p for m in mods for p in m.parameters()]
    opt = torch.optim.AdamW(params, lr=5e-5)
    lossf = nn.CrossEntropyLoss()
    ids = torch.randint(0, cfg.vocab_size, (a.batch, 512))
    labels = ids.clone()

    def step():
        with torch.autocast("cuda", dtype=torch.bfloat16):
            h = mods[0](ids.to(devs[0]))
            for m, d in zip(mods[1:-1], devs[1:-1]):
                h = m(h.to(d))
            logits = mods[-1](h.to(devs[-1]))
            loss = lossf(logits.view(-1, cfg.vocab_size).float(), labels.to(devs[-1]).view(-1))
        loss.backward()
        opt.step()
        opt.zero_grad(set_to_none=True)

    for _ in range(a.warmup):
        step()
    for d in range(n):
        torch.cuda.synchronize(d)
    t0 = time.perf_counter()
    for _ in range(a.steps):
        step()
    for d in range(n):
        torch.cuda.synchronize(d)
    ms = (time.perf_counter() - t0) / a.steps * 1e3
    print(json.dumps({"standin": "mp", "impl": "HF BERT modules, .to(device) hops, one process, bf16 autocast",
                      "model": f"bert-{a.model}-cased (untied head)", "n_gpus": n, "batch": a.batch, "ms_per_step": ms,
                      "samples_per_s": a.batch / ms * 1e3, "timing": "host clock around synchronised devices"}))


def run_pipe(a):
    from torch.distributed.pipelining import PipelineStage, ScheduleGPipe
    dist.init_process_group("nccl")
    rank, world = dist.get_rank(), dist.get_world_size()
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    cfg, mods = bert_modules(a.model)
    mods = [mods[0]] + [_Layer(m) for m in mods[1:-1]] + [mods[-1]]
    mine = np.array_split(np.arange(len(mods)), world)[rank]
    sub = nn.Sequential(*[mods[i] for i in mine]).to(dev).to(torch.bfloat16)
    stage = PipelineStage(sub, rank, world, dev)
    lossf = nn.CrossEntropyLoss()
    sched = ScheduleGPipe(stage, n_microbatches=a.chunks,
                          loss_fn=lambda out, tgt: lossf(out.view(-1, cfg.vocab_size).float(), tgt.view(-1)))
    opt = torch.optim.AdamW(sub.parameters(), lr=5e-5)
    ids = torch.randint(0, cfg.vocab_size, (a.batch, 512), device=dev)

    def step():
        if rank == 0:
            sched.step(ids)
        elif rank == world - 1:
            sched.step(target=ids)
        else:
            sched.step()
        opt.step()
        opt.zero_grad(set_to_none=True)

    ms = timed_loop(step, a.steps, a.warmup, world)
    if rank == 0:
        print(json.dumps({"standin": "pipe", "impl": f"torch.distributed.pipelining ScheduleGPipe, {a.chunks} micro-batches, bf16 weights",
                          "model": f"bert-{a.model}-cased (untied head)", "n_gpus": world, "batch": a.batch, "ms_per_step": ms,
                          "samples_per_s": a.batch / ms * 1e3}))
    dist.destroy_process_group()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("what", choices=["zero", "mp", "pipe"])
    ap.add_argument("--stage", type=int, default=0)
    ap.add_argument("--devices", type=int, default=None)
    ap.add_argument("--model", default="base", choices=["base", "large"])
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--chunks", type=int, default=4)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    a = ap.parse_args()
    {"zero": run_zero, "mp": run_mp, "pipe": run_pipe}[a.what](a)
