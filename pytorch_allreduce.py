"""All-reduce demonstration and bandwidth sweep.

Default behaviour is the reference's demo: three ranks each hold a different three-element
integer tensor, print it, sum it across ranks with one all-reduce, and print the result —
``[15, 27, 39]`` everywhere (pytorch_allreduce.py:17-33).  The reduction runs as this repository's
P2P kernel over NVLink peer memory on GPUs (NCCL only bootstraps), or over gloo on CPU.

``--sweep`` times all-reduce from 1 KB to 1 GB for the custom kernels and for NCCL and prints bus
bandwidth ``2(n-1)/n * bytes / time`` in GB/s per size (BASELINE.json config 2).  Times are CUDA
events on the launching stream, maximum over ranks.

    python pytorch_allreduce.py                       # the 3-rank demo
    python pytorch_allreduce.py --world-size 8 --sweep
"""
import argparse
import json
import os

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def create_process_group(rank, world_size, backend, port):
    from dtds_b200 import comm
    comm.init_process_group(backend=backend, rank=rank, world_size=world_size,
                            master_addr="127.0.0.1", master_port=port)


def all_reduce_example(rank, world_size, backend, port):
    from dtds_b200 import comm
    if backend == "nccl":
        torch.cuda.set_device(rank)
    create_process_group(rank, world_size, backend, port)
    device = torch.device("cuda", rank) if backend == "nccl" else torch.device("cpu")
    rows = {0: [1, 2, 3], 1: [10, 20, 30], 2: [4, 5, 6]}
    tensor = torch.tensor(rows.get(rank, [rank] * 3)).to(device)   # ranks > 2 contribute [r, r, r]
    print("Before AllReduce: Rank ", rank, " has data ", tensor)
    comm.all_reduce(tensor, op=dist.ReduceOp.SUM)
    print("After AllReduce:  Rank ", rank, " has data ", tensor)
    comm.destroy_process_group()


def _time_ms(fn, iters, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def sweep(rank, world_size, port, out_path, max_bytes):
    from dtds_b200 import comm
    torch.cuda.set_device(rank)
    create_process_group(rank, world_size, "nccl", port)
    ctx = comm.symm_context()
    buf = ctx.new_buffer(max_bytes)
    from dtds_b200.comm.symm import McastBuffer, all_reduce_nvls
    mc = McastBuffer(max_bytes)
    if rank == 0:
        print("NVLS multicast buffer:", "available" if mc.available else f"unavailable ({getattr(mc, 'error', 'no multicast address')})")
    results = []
    size = 1024
    while size <= max_bytes:
        n = size // 2
        sym = buf.tensor(torch.bfloat16, n)
        sym.fill_(1.0)
        plain = torch.ones(n, dtype=torch.bfloat16, device="cuda")
        out = torch.empty_like(plain)
        iters = 200 if size <= (1 << 20) else (50 if size <= (1 << 26) else 10)
        row = {"bytes": size}
        if size <= (512 << 10):
            row["custom_oneshot_ms"] = _time_ms(lambda: ctx.all_reduce(buf, 0, size, torch.bfloat16, out=out, algo=1), iters, 5)
        row["custom_twoshot_ms"] = _time_ms(lambda: ctx.all_reduce(buf, 0, size, torch.bfloat16, algo=2), iters, 5)
        if mc.available and size >= (1 << 16):
            mt = mc.tensor(torch.bfloat16, n)
            mt.fill_(1.0)
            all_reduce_nvls(ctx, mc, 0, size, torch.bfloat16, scale=1.0 / world_size)
            torch.cuda.synchronize()
            assert torch.all(mt == 1.0), "NVLS all-reduce produced a wrong result"
            row["custom_nvls_ms"] = _time_ms(lambda: all_reduce_nvls(ctx, mc, 0, size, torch.bfloat16), iters, 5)
        # what a user's `comm.all_reduce(tensor)` costs on an ordinary tensor: staged through the
        # symmetric scratch (copy in, kernel, copy out), reported next to the zero-copy numbers
        staged = torch.ones(n, dtype=torch.bfloat16, device="cuda")
        row["public_api_staged_ms"] = _time_ms(lambda: comm.all_reduce(staged), iters, 5)
        row["nccl_ms"] = _time_ms(lambda: dist.all_reduce(plain), iters, 5)
        best = min(v for k, v in row.items() if k.startswith("custom") and k.endswith("_ms"))
        f = 2 * (world_size - 1) / world_size * size / 1e6   # bus bytes / 1e6: ms -> GB/s
        row["custom_busbw_GBps"] = f / best
        row["public_api_busbw_GBps"] = f / row["public_api_staged_ms"]
        row["nccl_busbw_GBps"] = f / row["nccl_ms"]
        results.append(row)
        if rank == 0:
            algo = min((v, k) for k, v in row.items() if k.startswith("custom") and k.endswith("_ms"))[1][7:-3]
            row["custom_algo"] = algo
            print(f"{size:>12d} B  custom[{algo:>7s}] {row['custom_busbw_GBps']:9.2f} GB/s ({best * 1e3:9.1f} us)   "
                  f"comm.all_reduce(tensor) {row['public_api_busbw_GBps']:9.2f} GB/s   "
                  f"nccl {row['nccl_busbw_GBps']:9.2f} GB/s ({row['nccl_ms'] * 1e3:9.1f} us)", flush=True)
        size *= 4
    if rank == 0 and out_path:
        os.makedirs(os.path.dirname(out_path) or ".", exist_ok=True)
        with open(out_path, "w") as fh:
            json.dump({"world_size": world_size, "dtype": "bf16", "rows": results}, fh, indent=1)
    comm.destroy_process_group()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--world-size", type=int, default=3)
    ap.add_argument("--backend", choices=["nccl", "gloo"], default=None)
    ap.add_argument("--port", type=int, default=12355)
    ap.add_argument("--sweep", action="store_true")
    ap.add_argument("--max-bytes", type=int, default=1 << 30)
    ap.add_argument("--out", default="gpurun_out/allreduce_sweep.json")
    a = ap.parse_args()
    backend = a.backend or ("nccl" if torch.cuda.is_available() and torch.cuda.device_count() >= a.world_size else "gloo")
    if a.sweep:
        assert backend == "nccl", "the sweep needs one GPU per rank"
        mp.spawn(sweep, args=(a.world_size, a.port, a.out, a.max_bytes), nprocs=a.world_size)
    else:
        mp.spawn(all_reduce_example, args=(a.world_size, backend, a.port), nprocs=a.world_size)
