"""Put the unmodified reference under baseline/_ref (git-ignored; it travels with gpurun snapshots).

The prescribed install is tried first and its outcome recorded in baseline/_ref/INSTALL_LOG.txt:

    python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse \
        --target baseline/_ref /root/reference

The reference is a flat set of scripts with neither setup.py nor pyproject.toml, so pip refuses it;
the fallback is a byte-for-byte copy of the tree (sha256 of every file is logged so the copy can be
checked against /root/reference).  Nothing in the copy is edited: what the scripts need from the
environment (the removed transformers.AdamW, an offline bert-base-cased directory, an offline
wikitext directory) is supplied from outside by baseline/shims and baseline/make_assets.py.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DST = os.path.join(HERE, "_ref")
REF_SRC = os.environ.get("DTDS_REFERENCE_SRC", "/root/reference")


def _sha(path: str) -> str:
    h = hashlib.sha256()
    with open(path, "rb") as f:
        h.update(f.read())
    return h.hexdigest()


def installed() -> bool:
    return os.path.isfile(os.path.join(REF_DST, "data_parallel_training.py"))


def install(force: bool = False) -> str:
    """Returns a one-line outcome.  No-op when baseline/_ref is already populated."""
    if installed() and not force:
        return "present"
    if not os.path.isdir(REF_SRC):
        return f"source tree {REF_SRC} not found and baseline/_ref is empty"
    log = []
    cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--find-links",
           "/opt/wheelhouse", "--target", REF_DST, REF_SRC]
    try:
        p = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
        tail = (p.stdout + p.stderr).strip().splitlines()[-1:] or [""]
        log.append(f"$ {' '.join(cmd)}\nrc={p.returncode}: {tail[0]}")
        pip_ok = p.returncode == 0 and installed()
    except Exception as e:  # noqa: BLE001
        log.append(f"pip attempt raised {type(e).__name__}: {e}")
        pip_ok = False
    if not pip_ok:
        if os.path.isdir(REF_DST):
            shutil.rmtree(REF_DST)
        shutil.copytree(REF_SRC, REF_DST, ignore=shutil.ignore_patterns(".git", "__pycache__"))
        log.append(f"fallback: copied {REF_SRC} -> baseline/_ref unmodified")
        for root, _, files in sorted(os.walk(REF_DST)):
            for fn in sorted(files):
                p = os.path.join(root, fn)
                log.append(f"{_sha(p)}  {os.path.relpath(p, REF_DST)}")
    with open(os.path.join(REF_DST, "INSTALL_LOG.txt"), "w") as f:
        f.write("\n".join(log) + "\n")
    return "installed by pip" if pip_ok else "copied (pip: not installable, no setup.py/pyproject.toml)"


if __name__ == "__main__":
    print(install(force="--force" in sys.argv))
