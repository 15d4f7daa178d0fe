"""tcgen05 GEMM family vs an fp32 PyTorch reference (all operand layouts, epilogues, odd shapes)."""
import json
import os

import pytest
import torch

from dtds_b200.ops import kernels as K
from dtds_b200.ops import reference as ref

pytestmark = pytest.mark.gpu

OUT_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")


def _rand(*shape, scale=1.0):
    return (torch.randn(*shape, device="cuda", dtype=torch.float32) * scale).to(torch.bfloat16)


def _check(name, got, want, K_dim, report, out_is_bf16=True):
    got, want = got.float(), want.float()
    err = (got - want).abs()
    scale = want.abs().max().item() + 1e-6
    # bf16 output rounding (2^-8 relative) + fp32 accumulation-order noise
    tol = (2 ** -7 if out_is_bf16 else 2 ** -12) * scale + 1e-3
    bad = (err > tol + 2 ** -7 * want.abs()).sum().item() if out_is_bf16 else (err > tol).sum().item()
    entry = {"case": name, "max_err": err.max().item(), "ref_max": scale, "tol": tol,
             "bad": bad, "numel": got.numel(), "nan": bool(torch.isnan(got).any().item())}
    if bad:
        idx = (err > tol).nonzero()
        entry["first_bad"] = idx[:4].tolist()
        entry["bad_rows"] = sorted(set((idx[:, 0] // 32).tolist()))[:16]
        entry["bad_cols"] = sorted(set((idx[:, 1] // 32).tolist()))[:16]
    report.append(entry)
    return bad == 0 and not entry["nan"]


def _save(report, fname):
    os.makedirs(OUT_DIR, exist_ok=True)
    with open(os.path.join(OUT_DIR, fname), "w") as f:
        json.dump(report, f, indent=1)


SHAPES = [(128, 64, 64), (256, 256, 128), (2048, 768, 768), (2048, 2304, 768), (2048, 768, 3072),
          (512, 1024, 4096), (384, 200, 72), (130, 72, 200), (2048, 3072, 768)]


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
def test_gemm_layouts(a_mn, b_mn):
    torch.manual_seed(0)
    report, ok = [], True
    for (M, N, Kd) in SHAPES:
        for bn in (0, 64, 128, 256):
            if bn > 64 and N < bn:
                continue
            a = _rand(Kd, M) if a_mn else _rand(M, Kd)
            b = _rand(Kd, N) if b_mn else _rand(N, Kd)
            if (a.stride(0) % 8) or (b.stride(0) % 8):
                continue
            want = ref.gemm(a, b, a_mn_major=a_mn, b_mn_major=b_mn, out_dtype=torch.float32)
            got = K.gemm(a, b, a_mn_major=a_mn, b_mn_major=b_mn, block_n=bn)
            torch.cuda.synchronize()
            ok &= _check(f"M{M}N{N}K{Kd}_amn{int(a_mn)}_bmn{int(b_mn)}_bn{bn}", got, want, Kd, report)
    _save(report, f"gemm_layout_a{int(a_mn)}b{int(b_mn)}.json")
    assert ok, json.dumps([r for r in report if r["bad"] or r["nan"]][:6], indent=1)


def test_gemm_epilogues():
    torch.manual_seed(1)
    report, ok = [], True
    M, N, Kd = 512, 768, 256
    a, b = _rand(M, Kd), _rand(N, Kd, scale=0.1)
    bias = _rand(N)
    res = _rand(M, N)
    for act in (ref.ACT_NONE, ref.ACT_RELU, ref.ACT_GELU_ERF, ref.ACT_GELU_TANH):
        want, want_pre = ref.gemm(a, b, bias=bias, act=act, residual=res, out_dtype=torch.float32,
                                  want_pre=True)
        pre = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
        got = K.gemm(a, b, bias=bias, act=act, residual=res, pre_out=pre)
        torch.cuda.synchronize()
        ok &= _check(f"act{act}_out", got, want, Kd, report)
        ok &= _check(f"act{act}_pre", pre, want_pre, Kd, report)
        # backward-through-activation epilogue
        u = _rand(M, N)
        want = ref.gemm(a, b, act=act, dact_in=u, out_dtype=torch.float32)
        got = K.gemm(a, b, act=act, dact_in=u)
        ok &= _check(f"act{act}_dact", got, want, Kd, report)
    # fp32 output, accumulate, alpha
    acc = torch.randn(M, N, device="cuda")
    want = ref.gemm(a, b, alpha=0.5, out_dtype=torch.float32, accumulate_into=acc)
    got = K.gemm(a, b, alpha=0.5, out=acc.clone(), accumulate=True)
    ok &= _check("fp32_accumulate", got, want, Kd, report, out_is_bf16=False)
    accb = _rand(M, N)
    want = ref.gemm(a, b, out_dtype=torch.float32, accumulate_into=accb)
    got = K.gemm(a, b, out=accb.clone(), accumulate=True)
    ok &= _check("bf16_accumulate", got, want, Kd, report)
    _save(report, "gemm_epilogues.json")
    assert ok, json.dumps([r for r in report if r["bad"] or r["nan"]][:6], indent=1)


def test_gemm_vocab_padded_ld():
    """MLM decoder shape: N = 28996 (not a multiple of 8) written into a padded leading dimension,
    then consumed as a K-major / MN-major operand with K = 28996 (dgrad / wgrad of the decoder)."""
    torch.manual_seed(2)
    report, ok = [], True
    M, V, H, ld = 256, 28996, 768, 29056
    x, w = _rand(M, H), _rand(V, H, scale=0.05)
    bias = _rand(V)
    logits = torch.zeros(M, ld, dtype=torch.bfloat16, device="cuda")
    K.gemm(x, w, bias=bias, out=logits[:, :V])
    want = ref.gemm(x, w, bias=bias, out_dtype=torch.float32)
    ok &= _check("decoder_fwd", logits[:, :V], want, H, report)
    assert logits[:, V:].abs().max().item() == 0.0, "padding columns were written"
    d = (torch.randn(M, ld, device="cuda") * 0.1).to(torch.bfloat16)
    d[:, V:] = 0
    dv = d[:, :V]
    want = ref.gemm(dv, w, b_mn_major=True, out_dtype=torch.float32)  # dx = d · W
    got = K.gemm(dv, w, b_mn_major=True)
    ok &= _check("decoder_dgrad", got, want, V, report)
    want = ref.gemm(dv, x, a_mn_major=True, b_mn_major=True, out_dtype=torch.float32)  # dW = dᵀ·x
    got = K.gemm(dv, x, a_mn_major=True, b_mn_major=True)
    ok &= _check("decoder_wgrad", got, want, M, report)
    _save(report, "gemm_vocab.json")
    assert ok, json.dumps([r for r in report if r["bad"] or r["nan"]][:6], indent=1)


def test_gemm_throughput_report():
    """Times the BERT-base shapes and a large square GEMM; writes TFLOP/s to gpurun_out."""
    rows = []
    for (M, N, Kd) in [(2048, 2304, 768), (2048, 768, 768), (2048, 3072, 768), (2048, 768, 3072),
                       (2048, 28996, 768), (8192, 8192, 8192), (4096, 4096, 4096)]:
        a, b = _rand(M, Kd), _rand(N, Kd)
        ld = (N + 63) // 64 * 64
        out = torch.empty(M, ld, dtype=torch.bfloat16, device="cuda")[:, :N]
        for _ in range(3):
            K.gemm(a, b, out=out)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        iters = 10
        e0.record()
        for _ in range(iters):
            K.gemm(a, b, out=out)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
        for _ in range(3):
            torch.matmul(a, b.t(), out=c)
        t0.record()
        for _ in range(iters):
            torch.matmul(a, b.t(), out=c)
        t1.record()
        torch.cuda.synchronize()
        ms_ref = t0.elapsed_time(t1) / iters
        fl = 2.0 * M * N * Kd
        rows.append({"M": M, "N": N, "K": Kd, "ms": ms, "tflops": fl / ms / 1e9,
                     "cublas_ms": ms_ref, "cublas_tflops": fl / ms_ref / 1e9})
    _save(rows, "gemm_throughput.json")
    print(json.dumps(rows, indent=1))


# ------------------------------------------------------------------------------------ split-K
# csrc/gemm_splitk_sm100.cu was written after the round's GPU budget was spent: it compiles for
# sm_100a (96 registers, no spills) but has not run on hardware yet, and nothing uses it unless
# DTDS_SPLITK=1 or split_k is passed.  Opt in with DTDS_TEST_SPLITK=1.
_splitk = pytest.mark.skipif(os.environ.get("DTDS_TEST_SPLITK", "0") == "0",
                             reason="split-K GEMM not validated on hardware yet (set DTDS_TEST_SPLITK=1)")


@_splitk
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
def test_gemm_splitk_layouts(a_mn, b_mn):
    torch.manual_seed(0)
    report, ok = [], True
    for (M, N, Kd) in [(2048, 768, 3072), (2048, 768, 2304), (384, 200, 1032), (130, 136, 520), (768, 768, 2048)]:
        for bn, s in K.splitk_candidates(K._lib.lib(), torch.device("cuda"), M, N, Kd):
            a = _rand(Kd, M) if a_mn else _rand(M, Kd)
            b = _rand(Kd, N) if b_mn else _rand(N, Kd)
            want = ref.gemm(a, b, a_mn_major=a_mn, b_mn_major=b_mn, out_dtype=torch.float32)
            got = K.gemm(a, b, a_mn_major=a_mn, b_mn_major=b_mn, block_n=bn, split_k=s)
            ok &= _check(f"{M}x{N}x{Kd} bn{bn} s{s}", got, want, Kd, report)
            again = K.gemm(a, b, a_mn_major=a_mn, b_mn_major=b_mn, block_n=bn, split_k=s)   # counters were reset
            ok &= bool(torch.equal(got, again))   # fixed summation order: bit-reproducible
    _save(report, f"gemm_splitk_a{int(a_mn)}b{int(b_mn)}.json")
    assert report and ok, [r for r in report if r["bad"] or r["nan"]][:4]


@_splitk
def test_gemm_splitk_epilogues():
    torch.manual_seed(1)
    M, N, Kd, bn, s = 2048, 768, 3072, 256, 3
    report, ok = [], True
    x, w = _rand(M, Kd, scale=0.5), _rand(N, Kd, scale=0.05)
    bias, res = _rand(N), _rand(M, N)
    pre = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    got = K.gemm(x, w, bias=bias, act=ref.ACT_GELU_ERF, residual=res, pre_out=pre, block_n=bn, split_k=s)
    want, want_pre = ref.gemm(x, w, bias=bias, act=ref.ACT_GELU_ERF, residual=res, out_dtype=torch.float32,
                              want_pre=True)
    ok &= _check("bias+gelu+residual", got, want, Kd, report)
    ok &= _check("pre_out", pre, want_pre, Kd, report)
    # dgrad with the activation derivative fused (W consumed MN-major)
    dy, u = _rand(M, Kd, scale=0.5), _rand(M, N)
    wt = _rand(Kd, N, scale=0.05)
    got = K.gemm(dy, wt, b_mn_major=True, act=ref.ACT_GELU_ERF, dact_in=u, block_n=bn, split_k=s)
    want = ref.gemm(dy, wt, b_mn_major=True, act=ref.ACT_GELU_ERF, dact_in=u, out_dtype=torch.float32)
    ok &= _check("dgrad*gelu'", got, want, Kd, report)
    # wgrad accumulation into fp32
    a, b2 = _rand(2048, 768), _rand(2048, 768)
    acc = torch.randn(768, 768, device="cuda")
    want = acc + ref.gemm(a, b2, a_mn_major=True, b_mn_major=True, out_dtype=torch.float32)
    K.gemm(a, b2, a_mn_major=True, b_mn_major=True, out=acc, accumulate=True, block_n=128, split_k=4)
    ok &= _check("wgrad fp32 accumulate", acc, want, 2048, report, out_is_bf16=False)
    _save(report, "gemm_splitk_epilogues.json")
    assert ok, [r for r in report if r["bad"] or r["nan"]][:4]


@_splitk
def test_gemm_splitk_autotune_picks_a_valid_config(monkeypatch):
    monkeypatch.setattr(K, "_SPLITK", True)
    K._splitk_cache.clear()
    x, w = _rand(2048, 3072), _rand(768, 3072, scale=0.05)
    got = K.gemm(x, w)
    want = ref.gemm(x, w, out_dtype=torch.float32)
    report = []
    assert _check("autotuned", got, want, 3072, report), report
    (bn, s), = [v for k, v in K._splitk_cache.items() if k[:3] == (2048, 768, 3072)]
    assert bn in (64, 128, 256) and s >= 1


# ------------------------------------------------------------------------------ wide epilogue
_wide = pytest.mark.skipif(os.environ.get("DTDS_TEST_GEMM_WIDE", "0") == "0",
                           reason="16-warp GEMM epilogue not validated on hardware yet (set DTDS_TEST_GEMM_WIDE=1)")


@_wide
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
def test_gemm_wide_epilogue_layouts(a_mn, b_mn):
    """Same kernel with 16 epilogue warps (template parameter): all layouts, ragged shapes, fused
    epilogues, multi-tile CTAs (the TMEM double buffer hand-off counts 16 arrivals)."""
    torch.manual_seed(0)
    report, ok = [], True
    old = K.set_wide_epilogue(True)
    try:
        for (M, N, Kd) in [(2048, 768, 768), (2048, 3072, 768), (4096, 4096, 512), (392, 328, 200), (130, 136, 72)]:
            for bn in (128, 256):
                if N < bn:
                    continue
                a = _rand(Kd, M) if a_mn else _rand(M, Kd)
                b = _rand(Kd, N) if b_mn else _rand(N, Kd)
                want = ref.gemm(a, b, a_mn_major=a_mn, b_mn_major=b_mn, out_dtype=torch.float32)
                got = K.gemm(a, b, a_mn_major=a_mn, b_mn_major=b_mn, block_n=bn)
                ok &= _check(f"{M}x{N}x{Kd} bn{bn}", got, want, Kd, report)
        if not a_mn and not b_mn:
            x, w, bias, res = _rand(2048, 768, scale=0.5), _rand(3072, 768, scale=0.05), _rand(3072), _rand(2048, 3072)
            pre = torch.empty(2048, 3072, dtype=torch.bfloat16, device="cuda")
            got = K.gemm(x, w, bias=bias, act=ref.ACT_GELU_ERF, residual=res, pre_out=pre, block_n=256)
            want, want_pre = ref.gemm(x, w, bias=bias, act=ref.ACT_GELU_ERF, residual=res, out_dtype=torch.float32,
                                      want_pre=True)
            ok &= _check("bias+gelu+residual", got, want, 768, report) and _check("pre_out", pre, want_pre, 768, report)
    finally:
        K.set_wide_epilogue(old)
    _save(report, f"gemm_wide_a{int(a_mn)}b{int(b_mn)}.json")
    assert report and ok, [r for r in report if r["bad"] or r["nan"]][:4]
