"""One tiny forward + backward + optimizer step of the flagship model (BERT MLM) on cuda:0,
through the native sm_100a kernels.  Called by ``__graft_entry__.smoke()``."""
from __future__ import annotations

import torch


def run_smoke(verbose: bool = True) -> float:
    from .models import build_model, get_config
    from .ops import _lib
    from .optim import FusedAdam

    assert torch.cuda.is_available()
    _lib.lib()  # fail loudly if the native library is missing
    torch.cuda.set_device(0)
    torch.manual_seed(0)
    cfg = get_config("bert-tiny")
    model = build_model(cfg).to(device="cuda", dtype=torch.bfloat16)
    opt = FusedAdam(model.parameters(), lr=1e-3)
    ids = torch.randint(0, cfg.vocab_size, (2, 128), device="cuda")
    labels = ids.clone()
    labels[:, ::2] = -100
    _lib.reset_launch_count()
    losses = []
    for _ in range(3):
        opt.zero_grad()
        loss = model(ids, labels=labels).loss
        loss.backward()
        opt.step()
        losses.append(float(loss.item()))
    torch.cuda.synchronize()
    assert all(l == l for l in losses), f"NaN loss: {losses}"
    assert losses[-1] < losses[0], f"loss did not decrease: {losses}"
    if verbose:
        print(f"smoke ok: losses {['%.4f' % l for l in losses]}, native launches {_lib.launch_count()}")
    return losses[-1]
