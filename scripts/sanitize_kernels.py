"""One small launch of every single-GPU kernel family, meant to run under compute-sanitizer:

    compute-sanitizer --tool memcheck python scripts/sanitize_kernels.py
    compute-sanitizer --tool racecheck python scripts/sanitize_kernels.py   (shared-memory hazards)

(On the build pool compute-sanitizer is closed to tenants -- the wrapper answers "closed on this
pool" -- so there the script only runs plain, as a ragged-shape launch check: exit 0, 114 launches.)

Shapes are tiny and ragged on purpose (partial tiles, odd row counts) so out-of-bounds accesses in
tile tails show up.  Cross-GPU kernels are not covered: the sanitizer cannot see peer traffic; their
protocol is exercised by the stress tests in tests/test_gpu_multi.py instead.
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dtds_b200.ops import kernels as K  # noqa: E402
from dtds_b200.ops import _lib  # noqa: E402
from dtds_b200.ops.reference import ACT_GELU_ERF, ACT_RELU  # noqa: E402

dev = torch.device("cuda")
bf = torch.bfloat16
torch.manual_seed(0)


def rnd(*shape, scale=1.0):
    return (torch.randn(*shape, device=dev) * scale).to(bf)


# GEMM: every operand layout, ragged M/N/K (multiples of 8 only), each tile width, fused epilogues
for (M, N, Kd) in [(200, 136, 72), (128, 256, 64), (392, 328, 200)]:
    for a_mn in (False, True):
        for b_mn in (False, True):
            for bn in (64, 128, 256):
                a = rnd(Kd, M) if a_mn else rnd(M, Kd)
                b = rnd(Kd, N) if b_mn else rnd(N, Kd)
                K.gemm(a, b, a_mn_major=a_mn, b_mn_major=b_mn, block_n=bn)
x, w = rnd(200, 72), rnd(136, 72)
pre = torch.empty(200, 136, dtype=bf, device=dev)
K.gemm(x, w, bias=rnd(136), act=ACT_GELU_ERF, residual=rnd(200, 136), pre_out=pre)
K.gemm(rnd(200, 136), w, b_mn_major=True, act=ACT_GELU_ERF, dact_in=rnd(200, 72))
K.gemm(rnd(200, 136), rnd(200, 72), a_mn_major=True, b_mn_major=True,
       out=torch.zeros(136, 72, device=dev), accumulate=True)

# attention: every specialised mode and the generic one
B, S, h = 2, 256, 2
qkv, dout = rnd(B * S, 3 * h * 64, scale=0.5), rnd(B * S, h * 64)
kb = torch.zeros(B, S, device=dev)
kb[:, -17:] = -1e4
slopes = torch.tensor([0.5, 0.25], device=dev)
for causal, alibi, p, bias in [(False, None, 0.0, None), (False, None, 0.1, None), (False, None, 0.1, kb),
                               (True, None, 0.1, None), (True, slopes, 0.1, None), (True, slopes, 0.0, kb)]:
    kw = dict(scale=0.125, causal=causal, key_bias=bias, alibi=alibi, p_drop=p, seed=7)
    out, lse = K.attention_fwd(qkv, B, S, h, **kw)
    K.attention_bwd(qkv, out, dout, lse, B, S, h, **kw)

# normalisation / elementwise / reductions (column counts on every dispatch branch)
for rows, cols in [(130, 128), (77, 768), (33, 1024), (19, 2048)]:
    xx, res, g, bb = rnd(rows, cols), rnd(rows, cols), rnd(cols), rnd(cols)
    y, z, mean, rstd = K.layernorm_fwd(xx, res, g, bb, 1e-5, 0.1, 11)
    K.layernorm_bwd(rnd(rows, cols), z, mean, rstd, g, dz_extra=rnd(rows, cols), p_drop=0.1, seed=11,
                    want_dx_drop=True)
    K.colsum(xx)
    K.act_bwd(rnd(rows, cols), xx, ACT_GELU_ERF)
    K.act_fwd(xx, ACT_RELU)
    K.dropout(xx, 0.1, 5, res)
    K.softmax_rows_bwd(K.softmax_rows(xx), rnd(rows, cols))
    K.mse(xx, res)

# embedding, loss, optimizer
V, H = 1000, 128
ids = torch.randint(0, V, (3, 40), device=dev)
word, pos, typ = rnd(V, H), rnd(64, H), rnd(2, H)
e = K.embedding_fwd(ids, word, pos, typ, None, 2)
dword, dpos = torch.zeros(V, H, dtype=bf, device=dev), torch.zeros(64, H, dtype=bf, device=dev)
K.embedding_bwd(ids, rnd(3, 40, H), dword, dpos, pos_offset=2)
buf = rnd(120, 1008)
labels = torch.randint(0, V, (120,), device=dev)
labels[::5] = -100
K.cross_entropy(buf[:, :V], labels, vocab=V)
n = 10_007 * 8
p32, g16 = torch.randn(n, device=dev), rnd(n)
m, v, step = torch.zeros(n, device=dev), torch.zeros(n, device=dev), torch.zeros((), dtype=torch.int32, device=dev)
K.adam_flat(p32, g16, m, v, step, lr=1e-3, weight_decay=0.01, p_lp=torch.empty(n, dtype=bf, device=dev))
K.cast(p32, torch.empty(n, dtype=bf, device=dev))
torch.cuda.synchronize()
print("sanitize_kernels: done,", _lib.launch_count(), "native launches")
