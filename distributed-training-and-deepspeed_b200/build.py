"""Builds libdtds_b200.so (all sm_100a kernels + the symmetric-memory runtime) in-tree with nvcc.

The library has a plain C ABI (include/dtds_b200.h) and depends only on the CUDA runtime (linked
statically), so it needs no torch headers and compiles in seconds per file.  nvcc cross-compiles
for sm_100a without a GPU; the resulting .so travels to the GPU box with the snapshot.

    python -m dtds_b200.build            # incremental
    python -m dtds_b200.build --force    # rebuild everything
"""
from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
CSRC = PKG_DIR / "csrc"
INCLUDE = PKG_DIR.parent / "include"
LIB_DIR = PKG_DIR / "lib"
OBJ_DIR = LIB_DIR / "obj"
LIB_PATH = LIB_DIR / "libdtds_b200.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
    f"-I{INCLUDE}", f"-I{CSRC}",
]


def _nvcc() -> str:
    cand = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(cand):
        raise RuntimeError("nvcc not found; cannot build libdtds_b200.so")
    return cand


def sources() -> list[Path]:
    return sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cpp")))


def _digest(src: Path) -> str:
    h = hashlib.sha256()
    h.update(" ".join(NVCC_FLAGS).encode())
    h.update(src.read_bytes())
    for dep in sorted(list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h")) + list(INCLUDE.glob("*.h"))):
        h.update(dep.read_bytes())
    return h.hexdigest()


def _compile_one(src: Path, force: bool) -> tuple[Path, bool, str]:
    obj = OBJ_DIR / (src.stem + ".o")
    stamp = OBJ_DIR / (src.stem + ".sha")
    dig = _digest(src)
    if not force and obj.exists() and stamp.exists() and stamp.read_text() == dig:
        return obj, False, ""
    cmd = [_nvcc(), *NVCC_FLAGS, "-x", "cu", "-c", str(src), "-o", str(obj)]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{res.stdout}\n{res.stderr}")
    stamp.write_text(dig)
    return obj, True, res.stderr


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every source for sm_100a and link the shared library.  Returns its path."""
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    srcs = sources()
    objs, rebuilt, logs = [], False, []
    with cf.ThreadPoolExecutor(max_workers=min(8, len(srcs) or 1)) as ex:
        for obj, did, log in ex.map(lambda s: _compile_one(s, force), srcs):
            objs.append(obj)
            rebuilt |= did
            if did:
                logs.append(log)
    if verbose:
        for log in logs:
            print(log)
    if logs:
        (LIB_DIR / "ptxas.log").write_text("\n".join(logs))
    if rebuilt or not LIB_PATH.exists():
        cmd = [_nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a",
               "-o", str(LIB_PATH), *map(str, objs), "-lpthread", "-ldl"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    (LIB_DIR / "build.sha").write_text(_combined_digest())
    return LIB_PATH


def _combined_digest() -> str:
    h = hashlib.sha256()
    for src in sources():
        h.update(_digest(src).encode())
    return h.hexdigest()


def is_stale() -> bool:
    """True if the .so is missing or was built from different sources (lib/build.sha travels
    with the .so to the GPU box; the object files do not)."""
    stamp = LIB_DIR / "build.sha"
    return not (LIB_PATH.exists() and stamp.exists() and stamp.read_text() == _combined_digest())


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(path)
