#!/usr/bin/env python
"""Headline benchmark: BERT-base MLM data-parallel training throughput (samples/s).

Config = the reference's data_parallel_training.py defaults (BASELINE.json config 3): HF-shaped
BertForMaskedLM bert-base-cased, per-rank batch 4, sequence 512, AdamW lr 5e-5, DDP bucket 25 MiB,
bf16 compute with fp32 master weights, synthetic MLM token batches, random-init weights.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 5
    python bench.py --impl reference ...      # the unmodified reference script (baseline/_ref), fp32
    python bench.py --impl torch ...          # stand-in baseline: HF BERT + torch DDP + AdamW

Prints ONE JSON line (rank 0).  `value` is device-timed (CUDA events, max over ranks) over exactly
--steps replays of the captured training step; `e2e` re-measures the same steps through the
public API including the per-step pinned-host -> device input copy and device -> host loss read.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "torch"])
    ap.add_argument("--model", default="base", choices=["base", "large", "tiny"])
    ap.add_argument("--batch-size", type=int, default=4, help="per-rank batch (reference default 4)")
    ap.add_argument("--seq-len", type=int, default=512)
    ap.add_argument("--bucket-size", type=int, default=25)
    ap.add_argument("--no-graph", action="store_true", help="launch eagerly instead of replaying a CUDA graph")
    ap.add_argument("--no-overlap-optimizer", action="store_true",
                    help="one Adam launch after backward instead of per-bucket updates under backward")
    ap.add_argument("--adam-bucket-mb", type=float, default=32.0)
    ap.add_argument("--comm", default=None, choices=[None, "custom", "nccl"])
    return ap.parse_args()


class ClockSampler:
    """Samples SM clocks / throttle reasons with nvidia-smi while the timed region runs.  nvidia-smi
    needs ~0.1-0.3 s to deliver its first sample (longer on an 8-GPU box), which is as long as a
    20-step timed region, so the sampler is started before the warm-up steps and `stop(t0, t1)` keeps
    the samples whose timestamps fall into the timed window [t0, t1] (host clock, taken right
    after the synchronising barriers on both sides)."""

    Q = ("timestamp,index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.proc = gpu_index, None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20",
                 "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    @staticmethod
    def parse(out: str, t0: float | None = None, t1: float | None = None) -> dict:
        import datetime
        rows = []
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                ts = datetime.datetime.strptime(f[0], "%Y/%m/%d %H:%M:%S.%f").timestamp()
                rows.append((ts, float(f[2]), float(f[3]), [v.lower().startswith("active") for v in f[5:9]]))
            except ValueError:
                continue
        inside = [r for r in rows if t0 is not None and t1 is not None and t0 - 0.02 <= r[0] <= t1 + 0.02]
        use = inside or rows     # no sample inside a very short window: fall back to the whole run
        sm, mx = [r[1] for r in use], [r[2] for r in use]
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        reasons = sorted({n for r in use for n, on in zip(names, r[3]) if on})
        busy = [c for c in sm if c > 0.5 * max(mx or [1])] or sm
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(use), "samples_in_timed_window": len(inside), "reasons": reasons}

    def stop(self, t0: float | None = None, t1: float | None = None) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        return self.parse(out, t0, t1)


def reference_arm(args):
    """Runs the unmodified reference (baseline/_ref, see baseline/run_reference.py and DESIGN.md §4).
    Under torchrun only rank 0 starts it: the reference spawns its own N workers with mp.spawn."""
    if int(os.environ.get("RANK", "0")) != 0:
        return 0
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    try:
        import run_reference
        local = int(os.environ.get("LOCAL_RANK", "0"))
        out = run_reference.run(args.gpus, args.steps, args.warmup, batch=args.batch_size,
                                model="large" if args.model == "large" else "base", bucket_mb=args.bucket_size,
                                clock_sampler=ClockSampler(local))
    except Exception as e:  # noqa: BLE001
        import traceback
        out = {"impl": "reference", "unavailable": f"{type(e).__name__}: {e}; "
               + traceback.format_exc(limit=4).strip().replace("\n", " | ")[-800:]}
    print(json.dumps(out), flush=True)
    return 0


def make_batches(n, B, S, vocab, seed):
    """Synthetic MLM batches in pinned host memory: 15 % of positions carry a label, rest -100."""
    import torch
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(1000, vocab, (n, B, S), generator=g, dtype=torch.int64)
    labels = torch.full((n, B, S), -100, dtype=torch.int64)
    mask = torch.rand((n, B, S), generator=g) < 0.15
    labels[mask] = ids[mask]
    ids[mask] = 103  # [MASK]
    return ids.pin_memory(), labels.pin_memory()


def main():
    args = parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert world == args.gpus or world == 1 and args.gpus == 1, f"WORLD_SIZE {world} != --gpus {args.gpus}"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if args.comm:
        os.environ["DTDS_COMM"] = args.comm

    import dtds_b200  # noqa: F401
    from dtds_b200 import comm
    from dtds_b200.models import build_model, get_config
    from dtds_b200.ops import _lib

    if world > 1:
        comm.init_process_group(backend="nccl")
    model_name = {"base": "bert-base-cased", "large": "bert-large-cased", "tiny": "bert-tiny"}[args.model]
    cfg = get_config(model_name)
    B, S = args.batch_size, min(args.seq_len, cfg.max_position_embeddings)
    n_batches = args.steps + args.warmup
    ids_h, labels_h = make_batches(n_batches, B, S, cfg.vocab_size, seed=1234 + rank)
    torch.manual_seed(0)

    if args.impl == "torch":
        step_eager, step_graph, launches_per_step, describe = build_torch_baseline(args, cfg, dev, world, local)
    else:
        step_eager, step_graph, launches_per_step, describe = build_ours(args, cfg, dev, world, ids_h, labels_h)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- device-timed region: exactly K steps -----------------------------------
    ids_d, labels_d = ids_h.to(dev), labels_h.to(dev)
    sampler = ClockSampler(local)
    sampler.start()          # before the warm-up: nvidia-smi is delivering samples by the time the clock starts
    for i in range(args.warmup):
        step_graph(ids_d[i], labels_d[i])
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_start = time.time()
    e0.record()
    for i in range(args.steps):
        loss = step_graph(ids_d[args.warmup + i], labels_d[args.warmup + i])
    e1.record()
    barrier()
    t_end = time.time()
    clocks = sampler.stop(t_start, t_end)
    ms = e0.elapsed_time(e1)
    final_loss = float(loss.item())

    # ---------------- end-to-end region through the public API -------------------------------
    loss_host = torch.zeros((), dtype=torch.float32).pin_memory()
    barrier()
    t0 = time.perf_counter()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    for i in range(args.steps):
        loss = step_graph(ids_h[args.warmup + i], labels_h[args.warmup + i])   # H2D from pinned memory
        loss_host.copy_(loss.float(), non_blocking=True)                       # D2H of the step result
        torch.cuda.current_stream().synchronize()
    f1.record()
    barrier()
    e2e_wall_ms = (time.perf_counter() - t0) * 1e3
    e2e_ms = max(f0.elapsed_time(f1), e2e_wall_ms)

    evidence = describe.pop("_evidence", {})
    flat = describe.pop("_flat", None)
    if world > 1 and flat is not None:
        # every replica must hold bit-identical parameters after K + W + K steps
        bits = flat.param_flat.view(torch.int16).to(torch.int64)
        idx = torch.arange(bits.numel(), device=dev, dtype=torch.int64) % 8191 + 1
        cs = torch.stack([bits.sum(), (bits * idx).sum()])
        allcs = [torch.empty_like(cs) for _ in range(world)]
        dist.all_gather(allcs, cs)
        evidence["replicas_identical"] = bool(all(torch.equal(c, allcs[0]) for c in allcs))
        evidence["param_checksum"] = [int(x) for x in allcs[0].tolist()]
        ok = evidence["replicas_identical"]
        if "grad_vs_nccl_max_rel" in evidence:
            ok = ok and evidence["grad_vs_nccl_max_rel"] < 3e-2 and evidence["grad_vs_nccl_max_rel_embedding_bucket"] < 0.15
        evidence["multi_gpu_checks_passed"] = bool(ok)
        if not ok and rank == 0:   # reported, not fatal: the line below still carries the measurement
            print(f"[bench] multi-GPU correctness check failed: {evidence}", file=sys.stderr)

    t = torch.tensor([ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms = t.tolist()
    global_batch = B * world
    value = global_batch * args.steps / (ms / 1e3)
    e2e_value = global_batch * args.steps / (e2e_ms / 1e3)
    out = {
        "metric": "train samples/sec (DDP BERT MLM, data_parallel_training.py)",
        "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic MLM token batches (15% masked), random-init weights",
        "impl": args.impl,
        "config": {"model": cfg.name, "global_batch": global_batch, "per_gpu_batch": B, "seq_len": S,
                   "parallelism": f"dp{world}", "bucket_cap_mb": args.bucket_size, "optimizer": "AdamW lr=5e-5",
                   "l2": "per-step working set (params+grads+Adam state+activations, >2 GB) exceeds the 126 MB L2; "
                         "no explicit flush", **describe},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "samples/s", "ms_per_step": e2e_ms / args.steps,
                "h2d_bytes_per_step": 2 * B * S * 8, "d2h_bytes_per_step": 4},
        "gpu_launches": launches_per_step * args.steps,
        "final_loss": final_loss,
        **evidence,
    }
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def build_ours(args, cfg, dev, world, ids_h, labels_h):
    import torch
    from dtds_b200.models import build_model
    from dtds_b200.ops import _lib
    from dtds_b200.optim import HFAdamW
    from dtds_b200.parallel.ddp import DistributedDataParallel
    from dtds_b200.parallel.graph import GraphedStep

    model = build_model(cfg).to(device=dev, dtype=torch.bfloat16)
    model.train()
    net = DistributedDataParallel(model, device_ids=[dev.index], bucket_cap_mb=args.bucket_size) if world > 1 else model
    # Adam runs per bucket under backward (after the bucket's all-reduce when world > 1)
    overlap = not args.no_overlap_optimizer
    opt = HFAdamW(model.parameters(), lr=5e-5, overlap_with_backward=overlap and world == 1,
                  overlap_bucket_mb=args.adam_bucket_mb)

    def train_step(ids, labels):
        opt.zero_grad()
        loss = net(ids, labels=labels).loss
        loss.backward()
        opt.step()
        return loss

    ex = [ids_h[0].to(dev), labels_h[0].to(dev)]
    evidence = {}
    if world > 1:
        evidence.update(_check_bucket_against_nccl(net, ex, world))
    if overlap and world > 1:
        net.fuse_optimizer(opt)
    # count one eager step: launches of this library by C entry point, and library (NCCL) collectives
    n0, c0 = _lib.launch_count(), dict(_lib.call_counts)
    with _count_torch_collectives() as nccl_calls:
        train_step(*ex)
    per_step = _lib.launch_count() - n0
    calls = {k: v - c0.get(k, 0) for k, v in _lib.call_counts.items() if v - c0.get(k, 0)}
    evidence["comm_kernel_calls_per_step"] = {k: v for k, v in calls.items()
                                              if any(t in k for t in ("allreduce", "all_reduce", "reduce_scatter",
                                                                      "allgather", "all_gather", "zero_step", "broadcast"))}
    evidence["nccl_collective_calls_per_step"] = dict(nccl_calls)
    describe = {"native_kernels": True, "_evidence": evidence, "_flat": getattr(net, "flat", None)}
    if args.no_graph:
        return train_step, train_step, per_step, {"execution": "eager launches", **describe}
    gs = GraphedStep(train_step, ex, warmup=2)
    return train_step, gs, per_step, {"execution": "whole step replayed as one CUDA graph", **describe}


class _count_torch_collectives:
    """Counts torch.distributed collective calls (i.e. NCCL) made while active."""
    NAMES = ("all_reduce", "all_gather_into_tensor", "reduce_scatter_tensor", "broadcast", "all_gather", "reduce")

    def __enter__(self):
        import torch.distributed as dist
        self.counts, self._saved = {}, {}
        for n in self.NAMES:
            fn = getattr(dist, n, None)
            if fn is None:
                continue
            self._saved[n] = fn

            def wrapped(*a, _n=n, _fn=fn, **k):
                self.counts[_n] = self.counts.get(_n, 0) + 1
                return _fn(*a, **k)
            setattr(dist, n, wrapped)
        return self.counts

    def __exit__(self, *exc):
        import torch.distributed as dist
        for n, fn in self._saved.items():
            setattr(dist, n, fn)
        return False


def _check_bucket_against_nccl(net, ex, world):
    """One eager forward/backward without communication, NCCL all-reduce of a saved fp32 copy of the
    local gradients, then the same pass through the engine's own bucket kernels: the averaged
    gradient buffer must agree within bf16 rounding."""
    import torch
    import torch.distributed as dist
    from dtds_b200.ops import functional as F
    flat = net.flat
    F.manual_seed(4242)
    with net.no_sync():
        net(ex[0], labels=ex[1]).loss.backward()
    torch.cuda.synchronize()
    want = flat.grad_flat.float().clone()
    dist.all_reduce(want)
    want /= world
    F.manual_seed(4242)          # same dropout masks in the second pass
    net._accumulating = False    # drop the accumulated local gradients: start a fresh iteration
    net(ex[0], labels=ex[1]).loss.backward()
    torch.cuda.synchronize()
    got = flat.grad_flat.float()
    # Buckets that hold an embedding table are compared separately: its rows are summed with bf16
    # atomics (the [MASK] row collects ~300 contributions per step), whose order -- hence rounding --
    # differs between the two backward passes, independent of the all-reduce.
    worst = worst_atomic = 0.0
    for b in net.buckets:
        w, g = want[b.lo:b.hi], got[b.lo:b.hi]
        err = ((g - w).abs().max() / (w.abs().max() + 1e-12)).item()
        if any(getattr(p, "_grad_uses", 1) > 1 or p.dim() == 2 and p.shape[0] > 20000 for p in b.params):
            worst_atomic = max(worst_atomic, err)
        else:
            worst = max(worst, err)
    for p in flat.params:
        p.grad = None
    return {"grad_vs_nccl_max_rel": worst, "grad_vs_nccl_max_rel_embedding_bucket": worst_atomic,
            "grad_vs_nccl_buckets": len(net.buckets)}


def build_torch_baseline(args, cfg, dev, world, local):
    """Stand-in for the reference's stack on this image: HF BertForMaskedLM (random init from the
    in-repo config) + torch DistributedDataParallel over NCCL + torch.optim.AdamW, bf16 autocast."""
    import torch
    from transformers import BertConfig, BertForMaskedLM
    hf_cfg = BertConfig(vocab_size=cfg.vocab_size, hidden_size=cfg.hidden_size, num_hidden_layers=cfg.num_hidden_layers,
                        num_attention_heads=cfg.num_attention_heads, intermediate_size=cfg.intermediate_size,
                        max_position_embeddings=cfg.max_position_embeddings)
    model = BertForMaskedLM(hf_cfg).to(dev)
    model.train()
    net = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local], bucket_cap_mb=args.bucket_size) if world > 1 else model
    opt = torch.optim.AdamW(model.parameters(), lr=5e-5, eps=1e-6, weight_decay=0.0)

    def train_step(ids, labels):
        ids, labels = ids.to(dev, non_blocking=True), labels.to(dev, non_blocking=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            loss = net(input_ids=ids, labels=labels).loss
        loss.backward()
        opt.step()
        opt.zero_grad(set_to_none=True)
        return loss

    return train_step, train_step, 0, {"execution": "torch eager (HF BERT + torch DDP/NCCL + AdamW, bf16 autocast)",
                                       "native_kernels": False, "_evidence": {}, "_flat": None}


if __name__ == "__main__":
    sys.exit(main())
