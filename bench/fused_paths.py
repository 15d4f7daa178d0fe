"""Achieved fraction of roofline for the fused compute+collective kernels (run under torchrun).

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 bench/fused_paths.py

For each path the roofline time is the slower of (FLOPs / measured bf16 GEMM peak) and (bytes that
must cross NVLink per GPU / measured per-direction link bandwidth); both denominators come from
MEASURED_PEAKS.json / the profiling recipe (770 GB/s peer copy).  Times are CUDA events on the
launching stream, maximum over ranks, L2 flushed between iterations.  Writes
gpurun_out/fused_paths.json (rank 0).
"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dtds_b200 import comm  # noqa: E402
from dtds_b200.comm.symm import McastBuffer, all_reduce_nvls  # noqa: E402
from dtds_b200.ops import kernels as K  # noqa: E402

LINK_GBS = 770.0
try:
    PEAK_TFLOPS = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"]
except Exception:  # noqa: BLE001
    PEAK_TFLOPS = 1590.0


def timed(fn, iters=10, warmup=3, flush=None):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        dist.barrier()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    t = torch.tensor([sorted(ts)[len(ts) // 2]], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    global _last_stream_us
    _last_stream_us = timed_stream(fn)
    return t.item() * 1e3  # us


_last_stream_us = None


def timed_stream(fn, n=16):
    """Per-launch time of `n` launches issued back to back (no host barrier between them, so the
    host-side launch skew between ranks that a single timed launch waits out in its entry barrier is
    paid once, not per launch; operands may stay L2-resident).  Max over ranks."""
    torch.cuda.synchronize()
    dist.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    fn()
    s.record()
    for _ in range(n):
        fn()
    e.record()
    torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / n], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item() * 1e3


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    comm.init_process_group(backend="nccl")
    ctx = comm.symm_context()
    dev = torch.device("cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    rows = []

    def add(name, us, flops, link_bytes, note="", bb=None):
        t_c = flops / (PEAK_TFLOPS * 1e12) * 1e6
        t_l = link_bytes / (LINK_GBS * 1e9) * 1e6
        roof = max(t_c, t_l)
        bb = bb if bb is not None else _last_stream_us
        rows.append({"path": name, "time_us": round(us, 1), "roofline_us": round(roof, 1),
                     "fraction_of_roofline": round(roof / us, 3),
                     "time_us_back_to_back": round(bb, 1) if bb else None,
                     "fraction_of_roofline_back_to_back": round(roof / bb, 3) if bb else None,
                     "bound": "compute" if t_c > t_l else "nvlink",
                     "flops": flops, "nvlink_bytes_per_gpu_per_direction": link_bytes, "note": note})

    # ---- DDP gradient bucket (25 MiB bf16), in place, scale fused ------------------------------
    nbytes = 25 << 20
    buf = ctx.new_buffer(nbytes)
    buf.tensor(torch.bfloat16).fill_(1.0)
    us = timed(lambda: ctx.all_reduce(buf, 0, nbytes, torch.bfloat16, scale=1.0 / world, algo=2), flush=flush)
    add("ddp bucket all-reduce 25 MiB (two-shot P2P, ÷world + bf16 cast fused)", us, 0,
        (world - 1) / world * nbytes, "each rank pulls (n-1)/n and pushes (n-1)/n of the bucket")
    mc = McastBuffer(nbytes)
    if mc.available:
        mc.tensor(torch.bfloat16).fill_(1.0)
        us = timed(lambda: all_reduce_nvls(ctx, mc, 0, nbytes, torch.bfloat16, scale=1.0 / world), flush=flush)
        add("ddp bucket all-reduce 25 MiB (NVLS multimem)", us, 0, nbytes,
            "in-switch reduce: (n-1)/n out + 1/n in for the reduce, 1/n out + (n-1)/n in for the broadcast")

    # ---- ZeRO bucket: reduce-scatter + Adam + all-gather in one kernel (5e6 elements) ----------
    n = 5_000_000 // (8 * world) * (8 * world)
    per = n // world
    g, p = ctx.new_buffer(n * 2), ctx.new_buffer(n * 2)
    g.tensor(torch.bfloat16).normal_()
    master = torch.zeros(per, device=dev)
    m, v = torch.zeros(per, device=dev), torch.zeros(per, device=dev)
    step = torch.zeros((), dtype=torch.int32, device=dev)
    us = timed(lambda: ctx.zero_step(g, p, goff_elems=0, poff_elems=0, per_rank_elems=per, master=master, m=m,
                                     v=v, step=step, lr=1e-4, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0,
                                     grad_scale=1.0 / world), flush=flush)
    add("zero bucket 5e6 elts: reduce-scatter + Adam + all-gather (one kernel)", us, 0,
        (world - 1) / world * n * 2, "gradient slice pulled from n-1 peers, updated bf16 slice pushed to n-1 peers")

    gm, pm = McastBuffer(n * 2), McastBuffer(n * 2)
    if gm.available and pm.available:
        gm.tensor(torch.bfloat16, n).normal_()
        us = timed(lambda: ctx.zero_step(gm, pm, goff_elems=0, poff_elems=0, per_rank_elems=per, master=master, m=m,
                                         v=v, step=step, lr=1e-4, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0,
                                         grad_scale=1.0 / world), flush=flush)
        add("zero bucket 5e6 elts: NVLS reduce-scatter + Adam + NVLS all-gather (one kernel)", us, 0, n * 2 / world * (world - 1),
            f"multimem.ld_reduce of the owned slice, multimem.st of the updated slice ({gm.runtime} multicast runtime)")
        from dtds_b200.comm.symm import all_gather_nvls, reduce_scatter_nvls
        out = torch.empty(per, dtype=torch.bfloat16, device=dev)
        us = timed(lambda: reduce_scatter_nvls(ctx, gm, 0, per * 2, torch.bfloat16, out, scale=1.0 / world), flush=flush)
        add("reduce-scatter 10 MB bf16 (NVLS multimem.ld_reduce)", us, 0, (world - 1) / world * n * 2)
        us = timed(lambda: all_gather_nvls(ctx, pm, 0, per * 2, out), flush=flush)
        add("all-gather 10 MB bf16 (NVLS multimem.st)", us, 0, (world - 1) / world * n * 2)
    out = torch.empty(per, dtype=torch.bfloat16, device=dev)
    us = timed(lambda: ctx.reduce_scatter(g, 0, per * 2, torch.bfloat16, out, scale=1.0 / world), flush=flush)
    add("reduce-scatter 10 MB bf16 (P2P pull)", us, 0, (world - 1) / world * n * 2)
    us = timed(lambda: ctx.all_gather(p, 0, per * 2, out), flush=flush)
    add("all-gather 10 MB bf16 (P2P push)", us, 0, (world - 1) / world * n * 2)

    # ---- ZeRO-3: all-gather fused GEMM (bloom-560m shapes, batch 1 x 512 tokens per rank) -------
    for (M, N, Kd, tag) in [(512, 3072, 1024, "bloom-560m qkv"), (512, 4096, 1024, "bloom-560m h_to_4h"),
                            (2048, 12288, 4096, "6.7B-class qkv, 2048 tokens")]:
        per_w = N * Kd // world
        sb = ctx.new_buffer(per_w * 2)
        sb.tensor(torch.bfloat16).normal_(0, 0.02)
        torch.cuda.synchronize()
        dist.barrier()
        x = torch.randn(M, Kd, device=dev).to(torch.bfloat16)
        w_local = torch.empty(N, Kd, dtype=torch.bfloat16, device=dev)
        spec = K.GatherSpec([(sb.ptrs[r], per_w * 2) for r in range(world)], N * Kd * 2, dev, 32, Kd * 2,
                            local_begin=rank * per_w * 2)
        out = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
        us = timed(lambda: K.gemm_allgather(x, w_local, spec, out=out), flush=flush)
        bb_fused = _last_stream_us
        us_plain = timed(lambda: K.gemm(x, w_local, out=out), flush=flush)
        bb_plain = _last_stream_us
        add(f"all-gather fused GEMM {tag} (M={M} N={N} K={Kd})", us, 2.0 * M * N * Kd,
            (world - 1) / world * N * Kd * 2,
            f"same GEMM with the weight already local: {us_plain:.1f} us ({bb_plain:.1f} us back to back)", bb=bb_fused)

    if rank == 0:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        out = {"world_size": world, "link_GBps_per_direction": LINK_GBS, "bf16_peak_TFLOPs": PEAK_TFLOPS, "rows": rows}
        with open(os.path.join(ROOT, "gpurun_out", "fused_paths.json"), "w") as f:
            json.dump(out, f, indent=1)
        for r in rows:
            print(f"{r['path']}: {r['time_us']} us, roofline {r['roofline_us']} us ({r['bound']}), "
                  f"fraction {r['fraction_of_roofline']}  [{r['note']}]")
    comm.destroy_process_group()


if __name__ == "__main__":
    main()
