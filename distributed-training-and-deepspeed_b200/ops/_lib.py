"""ctypes loader for libdtds_b200.so.

Signatures are parsed from ``include/dtds_b200.h`` so the Python side cannot drift from the C
ABI.  On a machine with a GPU the library is mandatory: ``lib()`` raises instead of silently
falling back to PyTorch ops (CPU-only machines use the reference implementations in
``ops/reference.py``, which is what the non-GPU tests exercise).
"""
from __future__ import annotations

import ctypes
import re
import threading
from pathlib import Path

import torch

_PKG = Path(__file__).resolve().parent.parent
_HEADER = _PKG.parent / "include" / "dtds_b200.h"
_LIB_PATH = _PKG / "lib" / "libdtds_b200.so"

_lock = threading.Lock()
_lib = None
_launches = 0  # number of kernel-launching C calls made through this module


class NativeLibraryError(RuntimeError):
    pass


_CTYPES = {
    "int": ctypes.c_int,
    "unsigned": ctypes.c_uint,
    "unsigned int": ctypes.c_uint,
    "float": ctypes.c_float,
    "double": ctypes.c_double,
    "size_t": ctypes.c_size_t,
    "long long": ctypes.c_longlong,
    "unsigned long long": ctypes.c_ulonglong,
    "cudaStream_t": ctypes.c_void_p,
}


def _parse_header(text: str):
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    text = re.sub(r"//[^\n]*", " ", text)
    protos = {}
    for m in re.finditer(r"\b(int|void|size_t|unsigned long long|long long)\s+(dtds_\w+)\s*\(([^)]*)\)\s*;", text):
        ret, name, args = m.group(1), m.group(2), m.group(3)
        argtypes = []
        for a in [x.strip() for x in args.split(",") if x.strip()]:
            if a == "void":
                continue
            if "*" in a:
                argtypes.append(ctypes.c_void_p)
                continue
            a = re.sub(r"\bconst\b", "", a).strip()
            toks = a.split()
            ty = " ".join(toks[:-1]) if len(toks) > 1 else toks[0]
            if ty not in _CTYPES:
                raise NativeLibraryError(f"unknown C type '{ty}' in prototype of {name}")
            argtypes.append(_CTYPES[ty])
        restype = {"int": ctypes.c_int, "void": None, "size_t": ctypes.c_size_t,
                   "unsigned long long": ctypes.c_ulonglong, "long long": ctypes.c_longlong}[ret]
        protos[name] = (restype, argtypes)
    # a declaration this parser skipped would be called with ctypes' default int conversions
    declared = set(re.findall(r"\b(dtds_\w+)\s*\(", text))
    missed = declared - set(protos)
    if missed:
        raise NativeLibraryError(f"could not parse the prototype(s) of {sorted(missed)} in the C header")
    return protos


def available() -> bool:
    """True if the native library can be used (built, and a CUDA device is present)."""
    return torch.cuda.is_available() and _LIB_PATH.exists()


def lib():
    """Returns the loaded library, building nothing: raises with instructions if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not _LIB_PATH.exists():
            raise NativeLibraryError(
                f"{_LIB_PATH} is missing. Build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc cross-compiles sm_100a without a GPU). The framework does not fall back to "
                "PyTorch ops on a GPU machine.")
        handle = ctypes.CDLL(str(_LIB_PATH), mode=ctypes.RTLD_GLOBAL)
        for name, (restype, argtypes) in _parse_header(_HEADER.read_text()).items():
            try:
                fn = getattr(handle, name)
            except AttributeError as e:  # header and library out of sync
                raise NativeLibraryError(f"{name} declared in {_HEADER.name} but not exported") from e
            fn.restype = restype
            fn.argtypes = argtypes
        _lib = handle
    return _lib


def stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def ptr(t) -> int | None:
    """Device address of a tensor (None passes NULL)."""
    return None if t is None else t.data_ptr()


# kernels launched per C call (default 1); used for the launch accounting bench.py reports
_KERNELS_PER_CALL = {"dtds_cross_entropy": 2, "dtds_layernorm_bwd": 3,
                     "dtds_attention_bwd": 1, "dtds_adam_flat": 2}


call_counts: dict = {}  # C entry point name -> number of calls (tests assert on fused paths)


def check(rc: int, what: str, kernels: int | None = None) -> None:
    global _launches
    _launches += _KERNELS_PER_CALL.get(what, 1) if kernels is None else kernels
    call_counts[what] = call_counts.get(what, 0) + 1
    if rc != 0:
        raise NativeLibraryError(f"{what} failed with code {rc} (see stderr for details)")


def launch_count() -> int:
    return _launches


def reset_launch_count() -> None:
    global _launches
    _launches = 0


def use_native(*tensors) -> bool:
    """Dispatch rule: CUDA tensors -> native kernels (mandatory), CPU tensors -> reference ops."""
    for t in tensors:
        if t is not None and t.is_cuda:
            return True
    return False
