"""Allocation-free column sum (``dtds_colsum_noalloc``) against the regular entry point.

Who needs it: callers that must produce a bias gradient at a point where they cannot obtain
device memory — while a CUDA graph is being captured, from inside a caching allocator's own
callbacks / out-of-memory handling, or from a C host that was only handed tensors.  The scratch
the kernel needs sits in a fixed ``__device__`` array inside the library instead.

The test drives both entry points through the C ABI with ctypes (no torch op allocates between
the calls that matter), twice in a row — the second call proves the shared scratch was left
clean by the first — and checks the size limit is enforced with an error instead of a launch.
This file sorts last on purpose: if it fails under ``-x`` every other test has already run.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _colsum_both(L, x, out_dtype):
    rows, cols = x.shape
    stream = torch.cuda.current_stream().cuda_stream
    ref_out = torch.zeros(cols, dtype=out_dtype, device="cuda")
    scratch = torch.zeros(cols + cols // 64 + 130, dtype=torch.float32, device="cuda")
    rc = L.dtds_colsum(x.data_ptr(), rows, cols, x.stride(0), scratch.data_ptr(), ref_out.data_ptr(),
                       int(out_dtype == torch.float32), 0, stream)
    assert rc == 0
    new_out = torch.zeros(cols, dtype=out_dtype, device="cuda")
    rc = L.dtds_colsum_noalloc(x.data_ptr(), rows, cols, x.stride(0), new_out.data_ptr(),
                               int(out_dtype == torch.float32), 0, stream)
    assert rc == 0
    torch.cuda.synchronize()
    return ref_out, new_out


@pytest.mark.parametrize("rows,cols,out_dtype", [(2048, 768, torch.bfloat16), (300, 3072, torch.float32),
                                                 (4096, 28996, torch.float32), (64, 70, torch.bfloat16)])
def test_noalloc_matches_regular_entry_point_twice(rows, cols, out_dtype):
    from dtds_b200.ops import _lib
    L = _lib.lib()
    torch.manual_seed(0)
    ld = (cols + 7) // 8 * 8
    for call in range(2):  # two calls in a row share the built-in scratch
        x = torch.randn(rows, ld, device="cuda").to(torch.bfloat16)[:, :cols]
        want, got = _colsum_both(L, x, out_dtype)
        # same criterion as test_colsum in tests/test_gpu_ops.py: |a - b| <= atol + rtol * |b|
        atol, rtol = (0.5, 2e-2) if out_dtype == torch.bfloat16 else (2e-2, 1e-3)
        for other in (want.float(), x.float().sum(0)):
            err = (got.float() - other).abs()
            assert (err <= atol + rtol * other.abs()).all(), (call, err.max().item())


def test_noalloc_accumulate_and_size_limit():
    from dtds_b200.ops import _lib
    L = _lib.lib()
    stream = torch.cuda.current_stream().cuda_stream
    x = torch.randn(512, 256, device="cuda").to(torch.bfloat16)
    out = torch.ones(256, dtype=torch.float32, device="cuda")
    assert L.dtds_colsum_noalloc(x.data_ptr(), 512, 256, 256, out.data_ptr(), 1, 1, stream) == 0
    torch.cuda.synchronize()
    assert (out - (x.float().sum(0) + 1)).abs().max().item() < 2e-2
    # one column more than the built-in scratch holds: refused, nothing launched, output untouched
    big_cols = 32768 + 8
    out2 = torch.full((big_cols,), 7.0, dtype=torch.float32, device="cuda")
    y = torch.zeros(8, big_cols, dtype=torch.bfloat16, device="cuda")
    rc = L.dtds_colsum_noalloc(y.data_ptr(), 8, big_cols, big_cols, out2.data_ptr(), 1, 0, stream)
    torch.cuda.synchronize()
    assert rc != 0
    assert torch.all(out2 == 7.0)
