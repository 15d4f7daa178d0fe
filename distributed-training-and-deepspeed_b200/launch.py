"""Launcher: one process per GPU with the torchrun / DeepSpeed env contract.

Stands in for the ``deepspeed`` pdsh launcher the reference uses (scripts/launch-multinode.sh:5,
scripts/worker-prereqs.sh).  One 1/2/4/8-GPU NVSwitch box is the case the kernels are built for:

    python -m dtds_b200.launch --nproc-per-node 8 zero_dp_training.py --stage 3

Several boxes: either start it once per node (``--nnodes N --node-rank i --master-addr HOST``), or
give it a DeepSpeed-style hostfile (``host slots=N``) together with ``--ssh`` and it starts itself
on every listed host over ssh, first host = rendezvous address:

    python -m dtds_b200.launch --hostfile hostfile --ssh zero_dp_training.py --stage 2

Every child gets RANK, LOCAL_RANK, WORLD_SIZE, LOCAL_WORLD_SIZE, MASTER_ADDR, MASTER_PORT and, like
the DeepSpeed launcher, a ``--local_rank=N`` argument (scripts swallow it with parse_known_args,
zero_dp_training.py:121).  Across boxes the engines use the NCCL collectives (peer-mapped memory
and NVLS exist inside one NVSwitch domain only; see comm.custom_kernels_active).  If a child dies
the siblings are terminated and the launcher exits with the child's code (the only failure handling
the reference has is mp.spawn's join, data_parallel_training.py:76-81).
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
           env_extra: dict | None = None, nnodes: int = 1, node_rank: int = 0) -> int:
    if nnodes > 1 and master_port is None:
        raise ValueError("multi-node launch needs an agreed --master-port")
    port = master_port or free_port()
    procs = []
    for r in range(nproc):
        env = dict(os.environ)
        env.update({"RANK": str(node_rank * nproc + r), "LOCAL_RANK": str(r), "WORLD_SIZE": str(nnodes * nproc),
                    "LOCAL_WORLD_SIZE": str(nproc), "NODE_RANK": str(node_rank),
                    "MASTER_ADDR": master_addr, "MASTER_PORT": str(port)})
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


def ssh_fanout(hosts: dict[str, int], argv_tail: list[str], master_port: int, module: bool, no_local_rank: bool,
               ssh: str = "ssh", cwd: str | None = None, python: str | None = None) -> int:
    """Starts this launcher on every host of the hostfile (same working directory, same
    interpreter path) and waits for all of them; a failing node terminates the others."""
    import shlex
    names = list(hosts)
    slots = set(hosts.values())
    if len(slots) != 1:
        raise ValueError(f"hostfile slots differ across hosts ({hosts}); ranks are node_rank * slots + local_rank")
    nproc = slots.pop()
    cwd = cwd or os.getcwd()
    python = python or sys.executable
    procs = []
    for i, host in enumerate(names):
        inner = [python, "-m", "dtds_b200.launch", "--nnodes", str(len(names)), "--node-rank", str(i),
                 "--master-addr", names[0], "--master-port", str(master_port), "--nproc-per-node", str(nproc)]
        if module:
            inner.append("-m")
        if no_local_rank:
            inner.append("--no-local-rank")
        inner += argv_tail
        remote = f"cd {shlex.quote(cwd)} && " + " ".join(shlex.quote(a) for a in inner)
        procs.append(subprocess.Popen(shlex.split(ssh) + [host, remote]))
    code = 0
    alive = set(range(len(procs)))
    while alive:
        for i in list(alive):
            rc = procs[i].poll()
            if rc is None:
                continue
            alive.discard(i)
            if rc != 0 and code == 0:
                code = rc
                print(f"[launch] node {names[i]} exited with code {rc}; terminating the other nodes", file=sys.stderr)
                for j in alive:
                    procs[j].terminate()
        time.sleep(0.1)
    return code


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="dtds_b200.launch", description=__doc__.split("\n\n")[0])
    ap.add_argument("--nproc-per-node", "--num_gpus", dest="nproc", type=int, default=None)
    ap.add_argument("--master-addr", "--master_addr", dest="master_addr", default="127.0.0.1")
    ap.add_argument("--master-port", "--master_port", dest="master_port", type=int, default=None)
    ap.add_argument("--hostfile", default=None,
                    help="DeepSpeed-style hostfile (host slots=N): the local host's slot count, or every host with --ssh")
    ap.add_argument("--ssh", nargs="?", const="ssh", default=None, metavar="SSH_COMMAND",
                    help="with --hostfile: start one launcher per listed host over ssh")
    ap.add_argument("--nnodes", type=int, default=1)
    ap.add_argument("--node-rank", "--node_rank", dest="node_rank", type=int, default=0)
    ap.add_argument("--no-local-rank", action="store_true", help="do not append --local_rank=N")
    ap.add_argument("-m", "--module", action="store_true", help="run the target as a module")
    ap.add_argument("script")
    ap.add_argument("script_args", nargs=argparse.REMAINDER)
    a = ap.parse_args(argv)
    if a.ssh is not None:
        if not a.hostfile:
            ap.error("--ssh needs --hostfile")
        return ssh_fanout(parse_hostfile(a.hostfile), [a.script] + a.script_args, a.master_port or 29500,
                          a.module, a.no_local_rank, ssh=a.ssh)
    nproc = a.nproc
    if nproc is None and a.hostfile:
        slots = parse_hostfile(a.hostfile)
        nproc = slots.get(socket.gethostname(), next(iter(slots.values()), None))
    if nproc is None:
        nproc = default_nproc()
    if a.nnodes > 1 and a.master_port is None:
        a.master_port = 29500
    return launch(a.script, a.script_args, nproc, a.master_addr, a.master_port,
                  pass_local_rank=not a.no_local_rank, module=a.module, nnodes=a.nnodes, node_rank=a.node_rank)


if __name__ == "__main__":
    sys.exit(main())
