"""Reference arm of bench.py: run the UNMODIFIED reference DDP script and time K of its steps.

    baseline/_ref/data_parallel_training.py --batch-size B --training-steps (W+K)*N --device-count N
                                            --bucket-size MB --model base|large

is started as a child process from a scratch working directory that holds local `bert-base-cased/`
and `wikitext/` directories (baseline/make_assets.py), with baseline/shims first on PYTHONPATH
(restores transformers.AdamW, see shims/sitecustomize.py).  The reference shards `batch*steps` rows
over its N ranks, so `--training-steps (W+K)*N` makes every rank run W+K iterations of batch B: the
same per-GPU work as `bench.py --gpus N` (weak scaling).  The window between the end of iteration W
and the end of iteration W+K is timed inside the restored optimizer (device synchronised at both
ends, max over ranks); the reference's own `Total Training Time` lines are reported next to it.

The reference is fp32 end to end (no autocast) and its loop already contains the host->device copy of
each batch (DDP moves the DataLoader's CPU tensors) but never reads the loss back, so
`d2h_bytes_per_step` is 0 and `e2e` equals `value`: there is no device-only variant of its loop.
"""
from __future__ import annotations

import glob
import json
import os
import re
import subprocess
import sys
import tempfile
import time
import traceback

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
_TORCHRUN_VARS = ("RANK", "LOCAL_RANK", "WORLD_SIZE", "LOCAL_WORLD_SIZE", "GROUP_RANK", "GROUP_WORLD_SIZE",
                  "ROLE_RANK", "ROLE_WORLD_SIZE", "ROLE_NAME", "MASTER_ADDR", "MASTER_PORT", "OMP_NUM_THREADS")


def _tail(path: str, n: int = 25) -> str:
    try:
        with open(path, errors="replace") as f:
            lines = [l for l in f.read().replace("\r", "\n").splitlines() if l.strip() and "examples/s]" not in l]
        return "\n".join(lines[-n:])
    except OSError:
        return ""


def run(n_gpus: int, steps: int, warmup: int, batch: int = 4, model: str = "base", bucket_mb: int = 25,
        timeout_s: int = 1500, clock_sampler=None) -> dict:
    sys.path.insert(0, HERE)
    import install_ref
    import make_assets

    outcome = install_ref.install()
    if not install_ref.installed():
        return {"impl": "reference", "unavailable": f"baseline/_ref could not be populated: {outcome}"}
    script = os.path.join(install_ref.REF_DST, "data_parallel_training.py")
    work = os.path.join(tempfile.gettempdir(), f"dtds_ref_work_{os.getuid()}")
    model_name = "bert-large-cased" if model == "large" else "bert-base-cased"
    total_steps = (warmup + steps) * n_gpus
    try:
        make_assets.make_all(work, model_name, rows=max(256, batch * total_steps))
    except Exception:  # noqa: BLE001
        return {"impl": "reference", "unavailable": "offline assets could not be generated: "
                + traceback.format_exc(limit=3).strip().replace("\n", " | ")[-600:]}

    marks = os.path.join(work, f"marks_{os.getpid()}")
    for f in glob.glob(marks + ".rank*"):
        os.remove(f)
    env = {k: v for k, v in os.environ.items() if k not in _TORCHRUN_VARS and not k.startswith("TORCHELASTIC")}
    env.update({
        "PYTHONPATH": os.pathsep.join([os.path.join(HERE, "shims")] + [p for p in os.environ.get("PYTHONPATH", "").split(os.pathsep) if p]),
        "DTDS_REF_MARKS": f"{warmup},{steps},{marks}",
        "HF_DATASETS_OFFLINE": "1", "HF_HUB_OFFLINE": "1", "TRANSFORMERS_OFFLINE": "1",
        "HF_DATASETS_CACHE": os.path.join(work, ".hfcache"), "TOKENIZERS_PARALLELISM": "false",
    })
    cmd = [sys.executable, script, "--batch-size", str(batch), "--training-steps", str(total_steps),
           "--device-count", str(n_gpus), "--bucket-size", str(bucket_mb), "--model", model]
    log = os.path.join(work, f"reference_{os.getpid()}.log")
    if clock_sampler is not None:
        clock_sampler.start()
    t0 = time.time()
    with open(log, "w") as lf:
        try:
            rc = subprocess.run(cmd, cwd=work, env=env, stdout=lf, stderr=subprocess.STDOUT, timeout=timeout_s).returncode
        except subprocess.TimeoutExpired:
            rc = -9
    wall = time.time() - t0
    clocks = clock_sampler.stop() if clock_sampler is not None else None
    windows = []
    for f in sorted(glob.glob(marks + ".rank*")):
        with open(f) as fh:
            windows.append(float(fh.read().strip()))
        os.remove(f)
    text = _tail(log, 4000)
    totals = [float(x) for x in re.findall(r"Total Training Time: ([0-9.]+) seconds", text)]
    if rc != 0 or len(windows) != n_gpus:
        return {"impl": "reference", "unavailable":
                f"{' '.join(cmd[1:])} exited {rc} after {wall:.0f}s with {len(windows)}/{n_gpus} timing marks; log tail: "
                + _tail(log, 12).replace("\n", " | ")[-1500:]}
    sec = max(windows)
    value = batch * n_gpus * steps / sec
    return {
        "metric": "train samples/sec (DDP BERT MLM, data_parallel_training.py)",
        "value": value, "unit": "samples/s", "n_gpus": n_gpus, "steps": steps, "warmup": warmup,
        "ms_per_step": sec / steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "fp32", "data": "synthetic wikitext-shaped rows (local wikitext/ directory), random-init weights "
                                 "(local bert-*-cased/ directory); no network",
        "impl": "reference",
        "config": {"model": model_name, "global_batch": batch * n_gpus, "per_gpu_batch": batch, "seq_len": 512,
                   "parallelism": f"dp{n_gpus}", "bucket_cap_mb": bucket_mb, "optimizer": "transformers.AdamW lr=5e-5 "
                   "(restored as torch.optim.AdamW, eps 1e-6, wd 0)",
                   "execution": "unmodified reference script (mp.spawn, torch DDP over NCCL, HF BertForMaskedLM, fp32 eager)",
                   "l2": "per-step working set (fp32 params+grads+Adam state, >1.7 GB) exceeds the 126 MB L2; no explicit flush",
                   "same_config": True, "precision_note": "the reference has no autocast: fp32 (TF32 off), ours runs bf16",
                   "native_kernels": False, "reference_install": outcome},
        "clocks": clocks,
        "e2e": {"value": value, "unit": "samples/s", "ms_per_step": sec / steps * 1e3,
                "h2d_bytes_per_step": 2 * batch * 512 * 8, "d2h_bytes_per_step": 0,
                "note": "the reference loop copies each batch host->device itself and never reads the loss"},
        "gpu_launches": 0,
        "reference_total_training_time_s": totals, "reference_wall_s": wall,
        "timing": "window from end of iteration W to end of iteration W+K of the reference's own loop, device "
                  "synchronised at both ends, max over ranks (baseline/shims/sitecustomize.py)",
    }


if __name__ == "__main__":
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--model", default="base")
    a = ap.parse_args()
    print(json.dumps(run(a.gpus, a.steps, a.warmup, model=a.model)))
