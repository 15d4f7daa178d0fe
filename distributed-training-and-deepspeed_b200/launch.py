"""Single-box launcher: one process per GPU with the torchrun / DeepSpeed env contract.

Replaces the reference's cluster plumbing — the CDK stacks (aws/training_stack.py,
aws/multi_node_training_stack.py), the ssh key distribution and hostfile scripts
(scripts/generate-keys.sh, scripts/launch-multinode.sh) and the ``deepspeed`` pdsh launcher
(scripts/worker-prereqs.sh) — with a local spawner for one 1/2/4/8-GPU NVSwitch box:

    python -m dtds_b200.launch --nproc-per-node 8 zero_dp_training.py --stage 3

Every child gets RANK, LOCAL_RANK, WORLD_SIZE, MASTER_ADDR (127.0.0.1), MASTER_PORT and, like the
DeepSpeed launcher, a ``--local_rank=N`` argument (scripts swallow it with parse_known_args,
zero_dp_training.py:121).  A hostfile in DeepSpeed syntax (``host slots=N``) is accepted for
compatibility; only the local host's slot count is used.  If a child dies the siblings are
terminated and the launcher exits with the child's code (the only failure handling the reference
has is mp.spawn's join, data_parallel_training.py:76-81).
"""
from __future__ import annotations

import argparse
import os
import signal
import socket
import subprocess
import sys
import time


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def parse_hostfile(path: str) -> dict[str, int]:
    """``worker-1 slots=4`` lines -> {"worker-1": 4}."""
    slots = {}
    with open(path) as f:
        for line in f:
            line = line.split("#")[0].strip()
            if not line:
                continue
            host, *rest = line.split()
            n = 1
            for tok in rest:
                if tok.startswith("slots="):
                    n = int(tok.split("=", 1)[1])
            slots[host] = n
    return slots


def default_nproc() -> int:
    try:
        import torch
        return max(torch.cuda.device_count(), 1)
    except Exception:  # noqa: BLE001
        return 1


def launch(script: str, script_args: list[str], nproc: int, master_addr: str = "127.0.0.1",
           master_port: int | None = None, pass_local_rank: bool = True, module: bool = False,
           env_extra: dict | None = None) -> int:
    port = master_port or free_port()
    procs = []
    for r in range(nproc):
        env = dict(os.environ)
        env.update({"RANK": str(r), "LOCAL_RANK": str(r), "WORLD_SIZE": str(nproc),
                    "LOCAL_WORLD_SIZE": str(nproc), "MASTER_ADDR": master_addr, "MASTER_PORT": str(port)})
        if env_extra:
            env.update(env_extra)
        cmd = [sys.executable, "-u"] + (["-m", script] if module else [script])
        if pass_local_rank:
            cmd.append(f"--local_rank={r}")
        cmd += script_args
        procs.append(subprocess.Popen(cmd, env=env))
    code = 0
    try:
        alive = set(range(nproc))
        while alive:
            for r in list(alive):
                rc = procs[r].poll()
                if rc is None:
                    continue
                alive.discard(r)
                if rc != 0 and code == 0:
                    code = rc
                    print(f"[launch] rank {r} exited with code {rc}; terminating the other ranks", file=sys.stderr)
                    for q in alive:
                        procs[q].send_signal(signal.SIGTERM)
            time.sleep(0.05)
    except KeyboardInterrupt:
        for p in procs:
            p.send_signal(signal.SIGINT)
        code = 130
    finally:
        deadline = time.time() + 10
        for p in procs:
            try:
                p.wait(timeout=max(0.1, deadline - time.time()))
            except subprocess.TimeoutExpired:
                p.kill()
    return code


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="dtds_b200.launch", description=__doc__.split("\n\n")[0])
    ap.add_argument("--nproc-per-node", "--num_gpus", dest="nproc", type=int, default=None)
    ap.add_argument("--master-addr", "--master_addr", dest="master_addr", default="127.0.0.1")
    ap.add_argument("--master-port", "--master_port", dest="master_port", type=int, default=None)
    ap.add_argument("--hostfile", default=None, help="DeepSpeed-style hostfile; only local slots are used")
    ap.add_argument("--no-local-rank", action="store_true", help="do not append --local_rank=N")
    ap.add_argument("-m", "--module", action="store_true", help="run the target as a module")
    ap.add_argument("script")
    ap.add_argument("script_args", nargs=argparse.REMAINDER)
    a = ap.parse_args(argv)
    nproc = a.nproc
    if nproc is None and a.hostfile:
        slots = parse_hostfile(a.hostfile)
        nproc = slots.get(socket.gethostname(), next(iter(slots.values()), None))
    if nproc is None:
        nproc = default_nproc()
    return launch(a.script, a.script_args, nproc, a.master_addr, a.master_port,
                  pass_local_rank=not a.no_local_rank, module=a.module)


if __name__ == "__main__":
    sys.exit(main())
