"""Environment shim for running the UNMODIFIED reference scripts on this image (baseline/_ref).

Put on PYTHONPATH (only) for the reference arm of bench.py.  It supplies what the reference expects
from its 2023 environment and this image no longer has -- nothing of the reference is edited, none of
this repository's models / kernels / engines is imported:

* `transformers.AdamW` was removed in transformers 5.  It is restored as torch.optim.AdamW with the
  HF class's defaults (lr 1e-3, betas (0.9, 0.999), eps 1e-6, weight_decay 0.0, bias correction) --
  the library optimizer, multi-tensor path, which is faster than the per-tensor Python loop the
  reference really ran, so the stand-in favours the reference.
* `datasets` 4.6 trips over torchvision 0.26 (`torchvision.io.VideoReader` is gone) when it formats
  rows as torch tensors; it is told not to probe torchvision (text rows hold no video).
* Timing marks.  The reference only prints one wall-clock total that includes start-up (NCCL init,
  first-iteration autotuning), so a K-step window is cut out of its own loop: when
  DTDS_REF_MARKS="<warmup>,<steps>,<file>" is set, the restored AdamW synchronises the device and
  takes a timestamp at the end of its `warmup`-th and `(warmup+steps)`-th step() and each rank writes
  the elapsed seconds to <file>.rank<k>.  Two synchronisations per run, none inside the window.
* Any other `sitecustomize` further down sys.path (e.g. a harness hook) is chained afterwards.
"""
import importlib.abc
import importlib.util
import os
import sys
import time

_HERE = os.path.dirname(os.path.abspath(__file__))


def _make_adamw():
    import torch

    marks = os.environ.get("DTDS_REF_MARKS")
    warm = steps = None
    out = None
    if marks:
        w, k, out = marks.split(",", 2)
        warm, steps = int(w), int(k)

    class AdamW(torch.optim.AdamW):
        """transformers.AdamW (removed upstream) on top of torch.optim.AdamW, HF defaults."""

        def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.0,
                     correct_bias=True, no_deprecation_warning=True, **kw):
            super().__init__(params, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, **kw)
            self._n, self._t0 = 0, None

        def step(self, closure=None):
            r = super().step(closure)
            self._n += 1
            if warm is not None and self._n in (warm, warm + steps):
                torch.cuda.synchronize()
                now = time.perf_counter()
                if self._n == warm:
                    self._t0 = now
                else:
                    import torch.distributed as dist
                    rank = dist.get_rank() if dist.is_initialized() else 0
                    with open(f"{out}.rank{rank}", "w") as f:
                        f.write(f"{now - self._t0:.6f}\n")
            return r

    return AdamW


def _patch_transformers(mod):
    if "AdamW" not in vars(mod):
        setattr(mod, "AdamW", _make_adamw())


def _patch_datasets(mod):
    import datasets.config as c
    try:
        from torchvision.io import VideoReader  # noqa: F401
    except Exception:  # noqa: BLE001
        c.TORCHVISION_AVAILABLE = False


_PATCHES = {"transformers": _patch_transformers, "datasets": _patch_datasets}


class _PostImport(importlib.abc.MetaPathFinder):
    """Runs _PATCHES[name](module) right after `name` has been imported."""

    def __init__(self):
        self._busy = set()

    def find_spec(self, name, path=None, target=None):
        if name not in _PATCHES or name in self._busy:
            return None
        self._busy.add(name)
        try:
            spec = importlib.util.find_spec(name)
        finally:
            self._busy.discard(name)
        if spec is None or spec.loader is None:
            return None
        loader, orig_exec = spec.loader, spec.loader.exec_module

        def exec_module(module, _name=name):
            orig_exec(module)
            try:
                _PATCHES[_name](sys.modules.get(_name, module))
            except Exception as e:  # noqa: BLE001
                sys.stderr.write(f"[baseline shim] post-import patch of {_name} failed: {e}\n")

        loader.exec_module = exec_module
        return spec


sys.meta_path.insert(0, _PostImport())


def _chain_next_sitecustomize():
    for p in sys.path:
        if not p or os.path.abspath(p) == _HERE:
            continue
        f = os.path.join(p, "sitecustomize.py")
        if os.path.isfile(f):
            spec = importlib.util.spec_from_file_location("_chained_sitecustomize", f)
            mod = importlib.util.module_from_spec(spec)
            try:
                spec.loader.exec_module(mod)
            except Exception as e:  # noqa: BLE001
                sys.stderr.write(f"[baseline shim] chained sitecustomize {f} failed: {e}\n")
            return


_chain_next_sitecustomize()
