"""Which parameters differ between plain and overlapped Adam after one step (GPU debug aid)."""
import copy
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dtds_b200.models import build_model, get_config  # noqa: E402
from dtds_b200.optim import FusedAdam  # noqa: E402

torch.manual_seed(0)
cfg = get_config("bert-tiny").replace(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
a = build_model(cfg).to(device="cuda", dtype=torch.bfloat16)
b = copy.deepcopy(a)
ids = torch.randint(0, cfg.vocab_size, (4, 128), device="cuda")
labels = ids.clone()
labels[:, ::2] = -100
mb = float(os.environ.get("MB", "0.05"))
oa = FusedAdam(a.parameters(), lr=2e-3)
ob = FusedAdam(b.parameters(), lr=2e-3, overlap_with_backward=True, overlap_bucket_mb=mb)
w0 = {n: p.detach().float().clone() for n, p in a.named_parameters()}
for m, o in ((a, oa), (b, ob)):
    o.zero_grad()
    m(ids, labels=labels).loss.backward()
    o.step()
torch.cuda.synchronize()
ga, gb = oa.flats[0].grad_flat.float(), ob.flats[0].grad_flat.float()
print("grad flat max diff", (ga - gb).abs().max().item(), "buckets", len(ob._obuckets))
pb = dict(b.named_parameters())
for n, p in a.named_parameters():
    ua, ub = p.detach().float() - w0[n], pb[n].detach().float() - w0[n]
    bad = ((ua - ub).abs() > 1e-4).float().mean().item()
    if bad > 0:
        o = p._flat_offset
        print(f"{n:60s} off {o:8d} numel {p.numel():8d} differing {bad:.4f}  |ua| {ua.abs().mean():.2e} |ub| {ub.abs().mean():.2e}")
for bk in ob._obuckets[:40]:
    print("bucket", bk["lo"], bk["hi"], len(bk["params"]))
