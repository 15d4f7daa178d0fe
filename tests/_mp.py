"""Multi-process test harness: spawns `world` ranks running a function from tests/mp_workers.py
over gloo (CPU) or NCCL (one GPU per rank) with a 127.0.0.1 rendezvous on a free port."""
import os
import socket
import sys
import traceback

import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _entry(rank, world, backend, port, fn_name, args, errq):
    os.environ.update({"MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "RANK": str(rank),
                       "WORLD_SIZE": str(world), "LOCAL_RANK": str(rank)})
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    try:
        import torch.distributed as dist
        from dtds_b200 import comm
        if backend == "nccl":
            torch.cuda.set_device(rank)
        else:   # CPU ranks share the machine's cores instead of each spawning one thread per core
            torch.set_num_threads(max(1, (os.cpu_count() or 2) // world))
        comm.init_process_group(backend=backend, rank=rank, world_size=world)
        import mp_workers
        getattr(mp_workers, fn_name)(rank, world, *args)
        dist.barrier()
        dist.destroy_process_group()
    except Exception:  # noqa: BLE001
        errq.put((rank, traceback.format_exc()))
        raise


def run_distributed(fn_name: str, world: int = 2, backend: str = "gloo", args=(), timeout: int = 300):
    ctx = mp.get_context("spawn")
    errq = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_entry, args=(r, world, backend, port, fn_name, args, errq)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout)
    failed = []
    for r, p in enumerate(procs):
        if p.is_alive():
            p.kill()
            failed.append(f"rank {r} timed out")
        elif p.exitcode != 0:
            failed.append(f"rank {r} exit code {p.exitcode}")
    msgs = []
    while not errq.empty():
        msgs.append("rank %d:\n%s" % errq.get())
    assert not failed, "; ".join(failed) + "\n" + "\n".join(msgs)
