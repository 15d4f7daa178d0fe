"""Runs one multi-process worker from tests/mp_workers.py outside pytest (unbuffered output).
    python scripts/run_worker.py <worker> <world> <backend> [timeout_s] [args...]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
from _mp import run_distributed  # noqa: E402

if __name__ == "__main__":
    name, world, backend = sys.argv[1], int(sys.argv[2]), sys.argv[3]
    timeout = int(sys.argv[4]) if len(sys.argv) > 4 else 120
    def conv(a):
        try:
            return int(a)
        except ValueError:
            try:
                return float(a)
            except ValueError:
                return a
    args = tuple(conv(a) for a in sys.argv[5:])
    run_distributed(name, world=world, backend=backend, args=args, timeout=timeout)
    print(f"{name}: ok")
