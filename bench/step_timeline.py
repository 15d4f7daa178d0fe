"""Per-stream kernel timeline of one replayed training step (the bench.py step), via CUPTI through
torch.profiler -- for attribution only (which kernel owns the tail, what overlaps what); numbers
reported anywhere come from bench.py without a profiler.

    python bench/step_timeline.py                       # 1 GPU
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544 \
        bench/step_timeline.py --gpus 2

Rank 0 writes gpurun_out/step_timeline_N<gpus>.txt: every kernel of the middle profiled step with
stream, start and duration (us), then a summary (busy time per stream, tail after the last GEMM).
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench as B  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--no-graph", action="store_true")
    a = ap.parse_args()
    args = argparse.Namespace(gpus=a.gpus, steps=8, warmup=3, impl="ours", model="base", batch_size=4, seq_len=512,
                              bucket_size=25, no_graph=a.no_graph, no_overlap_optimizer=False, adam_bucket_mb=32.0,
                              comm=None)
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    from dtds_b200 import comm
    from dtds_b200.models import get_config
    import torch.distributed as dist
    if world > 1:
        comm.init_process_group(backend="nccl")
    cfg = get_config("bert-base-cased")
    ids_h, labels_h = B.make_batches(12, 4, 512, cfg.vocab_size, seed=1234 + rank)
    torch.manual_seed(0)
    _, step, _, _ = B.build_ours(args, cfg, dev, world, ids_h, labels_h)
    ids_d, labels_d = ids_h.to(dev), labels_h.to(dev)
    for i in range(4):
        step(ids_d[i], labels_d[i])
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for i in range(4, 9):
            step(ids_d[i], labels_d[i])
        torch.cuda.synchronize()
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and "emcpy" not in e.name.lower()
           and "memset" not in e.name.lower()]
    evs.sort(key=lambda e: e.time_range.start)
    if not evs:
        print("no CUDA kernel events recorded")
        return
    # split into steps at the embedding forward kernel (first kernel of a step)
    starts = [i for i, e in enumerate(evs) if "embedding_fwd" in e.name]
    if len(starts) >= 4:
        lo, hi = starts[2], starts[3]
    else:
        lo, hi = 0, len(evs)
    cur = evs[lo:hi]
    t0 = cur[0].time_range.start
    lines, busy, last_end = [], {}, {}
    for e in cur:
        s, d = e.time_range.start - t0, e.time_range.end - e.time_range.start
        stream = getattr(e, "device_resource_id", getattr(e, "device_index", 0))
        busy[stream] = busy.get(stream, 0) + d
        last_end[stream] = s + d
        lines.append(f"{s:9.1f} {d:8.1f}  s{stream:<3d} {e.name[:90]}")
    span = max(last_end.values())
    gemm_end = max((e.time_range.end - t0) for e in cur if "gemm" in e.name or "attn" in e.name)
    tail = [l for l, e in zip(lines, cur) if e.time_range.end - t0 > gemm_end]
    out = [f"step span {span:.1f} us, {len(cur)} kernels, world {world}; last GEMM/attention kernel ends at {gemm_end:.1f} us "
           f"(tail {span - gemm_end:.1f} us)", "busy us per stream: " + ", ".join(f"s{k}: {v:.0f}" for k, v in sorted(busy.items())),
           "kernels ending after the last GEMM/attention kernel:"] + tail + ["", "   start    dur(us) stream kernel"] + lines
    if rank == 0:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        path = os.path.join(ROOT, "gpurun_out", f"step_timeline_N{world}.txt")
        open(path, "w").write("\n".join(out) + "\n")
        print("\n".join(out[:3 + min(len(tail), 25)]))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
