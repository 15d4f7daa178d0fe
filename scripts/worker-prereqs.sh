#!/bin/bash
# Checks (and with --install tries to add) what a node needs before it can build and run this
# repository.  The reference installs pdsh + ninja here because its launcher fans out over pdsh and
# DeepSpeed JIT-compiles CUDA ops (scripts/worker-prereqs.sh:1); this launcher uses plain ssh and the
# library is built ahead of time with nvcc, so the list is: a driver, nvcc for sm_100a, Python deps.
set -u
fail=0
need() { command -v "$1" >/dev/null 2>&1 || { echo "missing: $1 ($2)"; fail=1; }; }
need nvidia-smi "NVIDIA driver"
need nvcc "CUDA 12.8+ toolkit, sm_100a support"
need python3 "Python 3.10+"
need ssh "multi-node launch"
if command -v nvcc >/dev/null 2>&1; then
  nvcc --list-gpu-arch 2>/dev/null | grep -q compute_100 || { echo "nvcc lacks compute_100a"; fail=1; }
fi
python3 - <<'PY' || fail=1
import importlib, sys
bad = [m for m in ("torch", "numpy", "tqdm", "tabulate") if importlib.util.find_spec(m) is None]
if bad:
    print("missing python packages:", ", ".join(bad)); sys.exit(1)
import torch
print("torch", torch.__version__, "cuda", torch.version.cuda, "gpus", torch.cuda.device_count())
PY
if [ "$fail" -ne 0 ] && [ "${1:-}" = "--install" ]; then
  python3 -m pip install -r "$(dirname "$0")/../requirements.txt"
  exec "$0"
fi
exit $fail
