"""End-to-end checks of the fused models on the GPU: bf16 native kernels vs the fp32 reference
path of the same modules, and whole-step CUDA-graph capture."""
import copy

import pytest
import torch

from dtds_b200.models import build_model, get_config
from dtds_b200.ops import _lib
from dtds_b200.optim import FusedAdam

pytestmark = pytest.mark.gpu


def _pair(name, **over):
    torch.manual_seed(0)
    cfg = get_config(name).replace(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, **over)
    ref = build_model(cfg).float()
    gpu = copy.deepcopy(ref).to(device="cuda", dtype=torch.bfloat16)
    with torch.no_grad():  # make the fp32 reference use exactly the bf16-rounded weights
        for pr, pg in zip(ref.parameters(), gpu.parameters()):
            pr.copy_(pg.float().cpu())
    return cfg, ref, gpu


@pytest.mark.parametrize("name", ["bert-tiny", "bloom-tiny", "opt-tiny"])
def test_model_matches_fp32_reference(name):
    cfg, ref, gpu = _pair(name)
    B, S = 2, 128
    ids = torch.randint(0, cfg.vocab_size, (B, S))
    labels = ids.clone()
    if cfg.arch == "bert":
        labels[:, ::3] = -100
    lr = ref(ids, labels=labels).loss
    lr.backward()
    n0 = _lib.launch_count()
    lg = gpu(ids.cuda(), labels=labels.cuda()).loss
    lg.backward()
    torch.cuda.synchronize()
    assert _lib.launch_count() > n0, "native kernels were not used"
    assert abs(lg.item() - lr.item()) < 3e-2 * max(1.0, abs(lr.item())), (lg.item(), lr.item())
    worst = 0.0
    for (n, pr), pg in zip(ref.named_parameters(), gpu.parameters()):
        assert pg.grad is not None, n
        gr, gg = pr.grad.float(), pg.grad.float().cpu()
        denom = gr.abs().max().item() + 1e-6
        rel = (gr - gg).abs().max().item() / denom
        worst = max(worst, rel)
        cos = torch.nn.functional.cosine_similarity(gr.flatten(), gg.flatten(), dim=0).item()
        # ReLU/GELU kinks turn bf16 rounding of pre-activations into a few percent of gradient noise
        # (the fp32-vs-bf16 CPU reference shows the same), so check direction tightly, magnitude loosely
        assert cos > 0.995 and rel < 0.25, f"{n}: rel grad err {rel:.3f} cos {cos:.4f} (ref max {denom:.3g})"
    print("worst relative grad error", worst)


def test_logits_path_and_padded_vocab():
    cfg, ref, gpu = _pair("bert-tiny", vocab_size=1003)  # not a multiple of 8 -> padded rows
    ids = torch.randint(0, cfg.vocab_size, (2, 128))
    lo_r = ref(ids).logits
    lo_g = gpu(ids.cuda()).logits
    assert lo_g.shape == lo_r.shape
    err = (lo_g.float().cpu() - lo_r).abs().max().item()
    assert err < 0.05 * lo_r.abs().max().item() + 0.05, err
    # loss outside the model, as model_parallel_training.py:73 does
    labels = ids.clone().cuda()
    from dtds_b200.ops import functional as Fn
    out = gpu(ids.cuda()).logits
    loss = Fn.cross_entropy(out, labels)
    loss.backward()
    ref_loss = torch.nn.functional.cross_entropy(lo_r.view(-1, cfg.vocab_size), ids.view(-1))
    assert abs(loss.item() - ref_loss.item()) < 0.05


def test_graphed_step_trains_and_matches_eager():
    from dtds_b200.parallel.graph import GraphedStep
    cfg, _, gpu = _pair("bert-tiny")
    eager = copy.deepcopy(gpu)
    B, S = 4, 128
    ids = torch.randint(0, cfg.vocab_size, (B, S), device="cuda")
    labels = ids.clone()
    labels[:, ::2] = -100

    def make_step(model, opt):
        def step(i, l):
            opt.zero_grad()
            loss = model(i, labels=l).loss
            loss.backward()
            opt.step()
            return loss
        return step

    opt_e = FusedAdam(eager.parameters(), lr=2e-3)
    step_e = make_step(eager, opt_e)
    ref_losses = [step_e(ids, labels).item() for _ in range(8)]

    opt_g = FusedAdam(gpu.parameters(), lr=2e-3)
    gs = GraphedStep(make_step(gpu, opt_g), [ids, labels], warmup=3)
    # warm-up (3) + capture (no execution) consumed optimizer steps: replays continue from step 3
    losses = [gs(ids, labels).item() for _ in range(5)]
    assert losses[-1] < losses[0]
    for a, b in zip(losses, ref_losses[3:]):
        assert abs(a - b) < 0.05 * max(1.0, abs(b)), (losses, ref_losses)


@pytest.mark.parametrize("graph", [False, True])
def test_adam_overlapped_with_backward_matches_plain(graph):
    """Per-bucket Adam on the optimizer's side stream under backward (optim/adam.py step_range)
    against one update after backward: same losses over 8 steps, eagerly and in a captured graph."""
    from dtds_b200.parallel.graph import GraphedStep
    cfg, _, a = _pair("bert-tiny")
    b = copy.deepcopy(a)
    ids = torch.randint(0, cfg.vocab_size, (4, 128), device="cuda")
    labels = ids.clone()
    labels[:, ::2] = -100

    def make_step(model, opt):
        def step(i, l):
            opt.zero_grad()
            loss = model(i, labels=l).loss
            loss.backward()
            opt.step()
            return loss
        return step

    opt_a = FusedAdam(a.parameters(), lr=2e-3)
    opt_b = FusedAdam(b.parameters(), lr=2e-3, overlap_with_backward=True, overlap_bucket_mb=0.05)
    assert len(opt_b._obuckets) > 5
    step_a, step_b = make_step(a, opt_a), make_step(b, opt_b)

    def flat(m):
        return torch.cat([p.detach().float().flatten() for p in m.parameters()])

    if graph:
        step_b = GraphedStep(step_b, [ids, labels], warmup=3)
        for _ in range(3):
            step_a(ids, labels)
    else:
        # first update: Adam's step-1 update is lr*sign(g), so run-to-run atomics noise flips some
        # elements; the noise floor is measured with a second plain run of the same model
        a2 = copy.deepcopy(a)
        step_a2 = make_step(a2, FusedAdam(a2.parameters(), lr=2e-3))
        w0 = flat(a)
        step_a(ids, labels), step_a2(ids, labels), step_b(ids, labels)
        ua, ua2, ub = flat(a) - w0, flat(a2) - w0, flat(b) - w0
        cos = torch.nn.functional.cosine_similarity
        floor = cos(ua, ua2, dim=0).item()
        got = cos(ua, ub, dim=0).item()
        assert got > 0.9 and got >= floor - 0.01, (got, floor)
    la = [step_a(ids, labels).item() for _ in range(8)]
    lb = [step_b(ids, labels).item() for _ in range(8)]
    assert la[-1] < la[0]
    for x, y in zip(la, lb):
        assert abs(x - y) < 0.02 * max(1.0, abs(x)), (la, lb)


def test_checkpoint_replays_fused_dropout():
    """Activation recomputation over fused blocks with dropout > 0: same gradients as without
    recomputation when both runs start from the same seed state (ops.functional.checkpoint)."""
    from dtds_b200.models import get_config
    from dtds_b200.models.bert import BertLayer
    from dtds_b200.ops import functional as Fn
    cfg = get_config("bert-tiny").replace(hidden_dropout_prob=0.2, attention_probs_dropout_prob=0.2)
    torch.manual_seed(0)
    layers = torch.nn.ModuleList([BertLayer(cfg) for _ in range(2)]).to(device="cuda", dtype=torch.bfloat16).train()
    x0 = torch.randn(2, 128, cfg.hidden_size, device="cuda").to(torch.bfloat16)

    def run(h):
        for m in layers:
            h = m(h)
        return h

    outs = {}
    for mode in ("plain", "checkpoint"):
        for p in layers.parameters():
            p.grad = None
        Fn.manual_seed(77)
        x = x0.clone().requires_grad_(True)
        y = run(x) if mode == "plain" else Fn.checkpoint(run, x)
        Fn.next_seed()                      # other consumers move the stream between forward and backward
        (y.float() ** 2).mean().backward()
        torch.cuda.synchronize()
        outs[mode] = [y.detach().float(), x.grad.float()] + [p.grad.float().clone() for p in layers.parameters()]
    for i, (a, b) in enumerate(zip(outs["plain"], outs["checkpoint"])):
        err = (a - b).abs().max().item() / (a.abs().max().item() + 1e-9)
        assert err < 2e-2, f"tensor {i}: recomputation changed the result by {err}"
