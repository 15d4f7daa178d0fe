"""The reference arm's environment (baseline/): the shim restores what the reference imports, the
offline assets resolve through the libraries' own local-directory lookup, and baseline/_ref is an
unmodified copy.  No GPU: the training loop itself is exercised on a B200 by `bench.py --impl reference`."""
import filecmp
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE = os.path.join(ROOT, "baseline")


def test_shim_restores_transformers_adamw_with_hf_defaults():
    code = ("import torch\nfrom transformers import AdamW\n"
            "o = AdamW([torch.nn.Parameter(torch.zeros(3))], lr=5e-5)\n"
            "assert issubclass(AdamW, torch.optim.AdamW)\n"
            "assert o.defaults['eps'] == 1e-6 and o.defaults['weight_decay'] == 0.0 and o.defaults['lr'] == 5e-5\n"
            "print('ok')\n")
    env = dict(os.environ, PYTHONPATH=os.path.join(BASE, "shims"))
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "ok" in r.stdout, r.stderr[-800:]


def test_reference_copy_is_unmodified():
    sys.path.insert(0, BASE)
    import install_ref
    if not os.path.isdir(install_ref.REF_SRC):
        pytest.skip("reference source tree not present on this machine")
    install_ref.install()
    assert install_ref.installed()
    for root, _, files in os.walk(install_ref.REF_SRC):
        if ".git" in root or "__pycache__" in root:
            continue
        for fn in files:
            src = os.path.join(root, fn)
            dst = os.path.join(install_ref.REF_DST, os.path.relpath(src, install_ref.REF_SRC))
            assert os.path.isfile(dst) and filecmp.cmp(src, dst, shallow=False), f"{dst} differs from the reference"


def test_offline_assets_resolve_like_the_hub_names(tmp_path):
    sys.path.insert(0, BASE)
    import make_assets
    make_assets.make_wikitext(str(tmp_path), rows=8)
    vocab = tmp_path / "bert-base-cased"
    vocab.mkdir()
    make_assets.make_vocab(str(vocab / "vocab.txt"))
    lines = (vocab / "vocab.txt").read_text().splitlines()
    assert len(lines) == 28996 and lines[0] == "[PAD]" and lines[100] == "[UNK]" and lines[101] == "[CLS]" \
        and lines[102] == "[SEP]" and lines[103] == "[MASK]"
    code = ("import datasets\n"
            "d = datasets.load_dataset('wikitext', 'wikitext-2-v1')['train']\n"
            "assert len(d) == 8 and len(d[0]['text'].split()) >= 512\nprint('ok')\n")
    env = dict(os.environ, PYTHONPATH=os.path.join(BASE, "shims"), HF_DATASETS_OFFLINE="1",
               HF_DATASETS_CACHE=str(tmp_path / ".hfcache"))
    r = subprocess.run([sys.executable, "-c", code], env=env, cwd=str(tmp_path), capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "ok" in r.stdout, r.stderr[-800:]
