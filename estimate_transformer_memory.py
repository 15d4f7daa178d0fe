"""Projected versus measured training memory of one transformer block, resized for a 180 GB B200.

The reference sizes a block at h=9216 / 72 heads / ffn 36864 / s=512 / b=4 in fp32 (17.0 GB
projected) for a 16 GB T4-class budget.  The closed forms are kept —

    model       = bytes_per_param * L * h * (13 + 12 h)
    gradients   = model
    activations = L * b * s * h * (67 + 9 a s / h) * bytes_per_value / 4        (arXiv:2205.05198)
    optimizer   = 8 bytes per parameter (Adam), 4 (SGD with momentum), else 0

— and evaluated for two configurations: the reference's (golden value 18,253,430,784 B ->
"17.0 GB") and a B200-sized one (default h=20480, 160 heads, s=2048, b=8: ~141 GB in fp32).  With
``--fused`` the quadratic attention term is dropped (the fused attention kernel never materialises
the [b, a, s, s] scores) and dropout masks cost nothing.

After the projection the tool measures the real thing on the available accelerator the way the
reference does: allocator deltas after moving the model, after forward (against the hook-counted
activations), after MSE backward and after ``Adam.step()``.  On a CPU-only machine it prints the
projections and the hook-counted estimates and stops.  (Quirk Q1 of the reference —
``project_transformer_memory`` reading a global ``model`` — is fixed: the parameter count comes
from the arguments.)
"""
import argparse
from types import SimpleNamespace

import torch

from estimate_nn_memory import ActivationCounter, get_model_memory, register_hooks_recursive
from model.transformer import TransformerBlock
from util import format_size, get_device


def get_current_memory_allocation(device):
    if device == "cuda":
        return torch.cuda.memory_allocated()
    if device == "mps":
        return torch.mps.current_allocated_memory()
    raise ValueError(f"Unsupported device: {device}")


def optimizer_bytes_per_param(optimizer) -> int:
    if isinstance(optimizer, str):
        name = optimizer.lower()
        return {"adam": 8, "adamw": 8, "sgd_momentum": 4, "sgd": 0}[name]
    if type(optimizer) is torch.optim.SGD:
        return 4 if any(g.get("momentum", 0) != 0 for g in optimizer.param_groups) else 0
    if type(optimizer) in (torch.optim.Adam, torch.optim.AdamW) or type(optimizer).__name__ == "FusedAdam":
        return 8
    raise ValueError(f"Unsupported optimizer: {optimizer}")


def get_optimizer_memory(model: torch.nn.Module, optimizer) -> int:
    return sum(p.numel() for p in model.parameters()) * optimizer_bytes_per_param(optimizer)


def transformer_param_count(layers, hidden_size):
    return layers * hidden_size * (13 + 12 * hidden_size)


def project_transformer_memory(layers, hidden_size, num_attention_heads, batch_size, sequence_length,
                               optimizer="adam", bytes_per_value=4, fused_attention=False, breakdown=False):
    params = transformer_param_count(layers, hidden_size)
    model_memory = bytes_per_value * params
    gradient_memory = model_memory
    quad = 0 if fused_attention else (9 * num_attention_heads * sequence_length) / hidden_size
    linear = 67 if not fused_attention else 67 - 3   # three dropout masks are never stored when fused
    activation_memory = layers * batch_size * sequence_length * hidden_size * (linear + quad) * bytes_per_value / 4
    optimizer_memory = params * optimizer_bytes_per_param(optimizer)
    total = model_memory + gradient_memory + activation_memory + optimizer_memory
    if breakdown:
        return {"model": model_memory, "gradients": gradient_memory, "activations": activation_memory,
                "optimizer": optimizer_memory, "total": total}
    return total


def block_config(hidden, heads, seq, dropout=0.1):
    return SimpleNamespace(hidden_size=hidden, num_attention_heads=heads, enable_bias=True,
                           max_position_embeddings=seq, ffn_dim=4 * hidden, dropout=dropout)


def _report(title, measured, estimated):
    print(f"Measured {title}: {format_size(measured)}")
    print(f"Estimated {title}: {format_size(estimated)}")
    print(f"Percent difference: {abs(measured - estimated) / estimated * 100:.2f}%")
    print("-" * 80)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--hidden", type=int, default=None)
    ap.add_argument("--heads", type=int, default=None)
    ap.add_argument("--seq", type=int, default=None)
    ap.add_argument("--batch", type=int, default=None)
    ap.add_argument("--reference-config", action="store_true", help="measure the T4-sized configuration")
    ap.add_argument("--dtype", choices=["fp32", "bf16"], default="fp32")
    ap.add_argument("--fused", action="store_true", help="fused attention path (needs bf16, head_dim 64)")
    ap.add_argument("--no-measure", action="store_true")
    a = ap.parse_args(argv)

    ref = dict(hidden=9216, heads=72, seq=512, batch=4)
    big = dict(hidden=20480, heads=160, seq=2048, batch=8)
    print(f"Projected total memory usage: "
          f"{format_size(project_transformer_memory(1, ref['hidden'], ref['heads'], ref['batch'], ref['seq']))}"
          f"  (reference configuration h={ref['hidden']} a={ref['heads']} s={ref['seq']} b={ref['batch']}, fp32 + Adam)")
    proj_big = project_transformer_memory(1, big["hidden"], big["heads"], big["batch"], big["seq"], breakdown=True)
    print(f"Projected total memory usage: {format_size(proj_big['total'])}"
          f"  (B200 configuration h={big['hidden']} a={big['heads']} s={big['seq']} b={big['batch']}, fp32 + Adam; "
          f"HBM capacity 180 GB)")
    chosen = dict(ref if a.reference_config else big)
    for k in ("hidden", "heads", "seq", "batch"):
        if getattr(a, k) is not None:
            chosen[k] = getattr(a, k)
    bpv = 4 if a.dtype == "fp32" else 2
    proj = project_transformer_memory(1, chosen["hidden"], chosen["heads"], chosen["batch"], chosen["seq"],
                                      bytes_per_value=bpv, fused_attention=a.fused, breakdown=True)
    print(f"Selected configuration {chosen} dtype={a.dtype} fused={a.fused}: "
          + ", ".join(f"{k} {format_size(v)}" for k, v in proj.items()))
    print("-" * 80)

    device = get_device()
    if a.no_measure or device == "cpu":
        print("No accelerator: skipping the measured half (allocator statistics need CUDA).")
        return proj
    dtype = torch.float32 if a.dtype == "fp32" else torch.bfloat16
    config = block_config(chosen["hidden"], chosen["heads"], chosen["seq"])
    model = TransformerBlock(config, fused=a.fused).to(dtype)
    base = get_current_memory_allocation(device)
    model.to(device)
    with_model = get_current_memory_allocation(device) - base
    estimated_model = get_model_memory(model)
    _report("Model Memory", with_model, estimated_model)

    counter = ActivationCounter(dropout_bytes_per_element=0 if dtype == torch.bfloat16 else 1)
    register_hooks_recursive(model, counter)
    inputs = torch.randn(chosen["batch"], chosen["seq"], chosen["hidden"], device=device, dtype=dtype)
    before = get_current_memory_allocation(device)
    outputs = model(inputs)
    counter.add_activations(inputs)
    forward_bytes = get_current_memory_allocation(device) - before + inputs.numel() * inputs.element_size()
    _report("Activation Memory", forward_bytes, counter.activation_bytes)

    labels = torch.randn_like(outputs)
    before = get_current_memory_allocation(device)
    loss = torch.nn.functional.mse_loss(outputs.float(), labels.float())
    loss.backward(retain_graph=True)
    grad_bytes = get_current_memory_allocation(device) - before
    _report("Gradient Memory", grad_bytes, estimated_model)

    optimizer = torch.optim.Adam(model.parameters(), lr=1e-3)
    before = get_current_memory_allocation(device)
    optimizer.step()
    opt_bytes = get_current_memory_allocation(device) - before
    _report("Optimizer Memory", opt_bytes, get_optimizer_memory(model, optimizer) * (1 if dtype == torch.float32 else 0.5))
    print(f"Actual total memory usage: {format_size(get_current_memory_allocation(device))}")
    return proj


if __name__ == "__main__":
    main()
