"""Autograd wrappers around the sm_100a kernels.

Each ``torch.autograd.Function`` here is one fused forward and one fused backward built from the
raw kernels in ``kernels.py``; CPU tensors run the same code through the fp32 reference ops, so
the engines and CLIs can be tested without a GPU.

Gradient emission protocol (used by the DDP / ZeRO engines, see parallel/flat.py): if a
parameter carries ``main_grad`` (a view into a flat gradient buffer) the weight-gradient GEMM
writes — or accumulates — straight into it, no ``.grad`` tensor is produced, and the parameter's
``_on_grad_ready`` callback fires as soon as the kernel is enqueued.  That is what lets the
bucketed all-reduce / reduce-scatter start while the rest of backward is still running, with no
pack or unpack copies ("DDP variant writes dW straight into the flat bucket").
"""
from __future__ import annotations

import os
import math

import torch

from . import kernels as K
from .reference import ACT_NONE

_SEED_START = 0x5EED
_seed_counter = _SEED_START
_base_seed = 0


def manual_seed(seed: int) -> None:
    """Seeds the dropout streams of the fused kernels (independent of torch's generator)."""
    global _seed_counter, _base_seed
    _base_seed = int(seed) & 0xFFFFFFFF
    _seed_counter = _SEED_START


def next_seed() -> int:
    global _seed_counter
    n = _seed_counter
    _seed_counter = n + 1
    return ((_base_seed << 32) ^ (n * 0x9E3779B1)) & 0x7FFFFFFFFFFFFFFF


def get_seed_state() -> tuple:
    """Position of the host-side dropout seed stream.  Code that re-runs a forward and needs the
    SAME masks (activation recomputation) saves it before the first run and restores it with
    `set_seed_state` before the re-run."""
    return (_base_seed, _seed_counter)


def set_seed_state(state: tuple) -> None:
    global _seed_counter, _base_seed
    _base_seed, _seed_counter = state


def checkpoint(fn, *args):
    """torch.utils.checkpoint (non-reentrant) for code built on the fused ops.  The fused kernels draw
    dropout masks from the host-side seed stream above, which torch's checkpoint neither saves nor
    restores; the backward nodes of the first run keep their seeds, so the recomputation must
    regenerate exactly the same masks: it replays the stream from the position the first run
    started at and then puts the stream back where it was."""
    from torch.utils.checkpoint import checkpoint as _ckpt
    start = get_seed_state()
    first = [True]

    def run(*a):
        if first[0]:
            first[0] = False
            return fn(*a)
        here = get_seed_state()
        set_seed_state(start)
        try:
            return fn(*a)
        finally:
            set_seed_state(here)

    # torch's own RNG is not used by the fused ops; reading it is also illegal under stream capture
    return _ckpt(run, *args, use_reentrant=False, preserve_rng_state=False)


# ------------------------------------------------------------------ gradient emission helpers
def grad_target(param):
    """(buffer, accumulate) for a parameter that owns a flat-buffer gradient view, else (None, _)."""
    mg = getattr(param, "main_grad", None)
    if mg is None:
        return None, False
    return mg, bool(getattr(param, "_grad_written", False))


def mark_shared_grad(param, uses: int) -> None:
    """Declares that `uses` backward producers write this parameter's gradient (tied embeddings).
    Models call it at the top of forward, not only in __init__: ``copy.deepcopy`` of a Parameter
    drops Python attributes, and a lost mark would fire the gradient-ready callback (bucket
    all-reduce / Adam) after the first producer."""
    if getattr(param, "_grad_uses", 1) != uses:
        from ..parallel.flat import grad_uses
        param._grad_uses = uses
        param._uses_pending = grad_uses(param)


def grad_done(param) -> None:
    param._grad_written = True
    pending = getattr(param, "_uses_pending", 1) - 1
    param._uses_pending = pending
    if pending <= 0:
        cb = getattr(param, "_on_grad_ready", None)
        if cb is not None:
            cb(param)


# ------------------------------------------------------------------ weight-gradient side stream
# dgrad feeds the next layer's backward, wgrad feeds nobody until the optimizer (or the bucket
# collective).  At the reference's batch sizes one GEMM has 50-300 tiles for 148 SMs, so running
# the wgrad (+ bias column sum) of a layer on a second stream lets it fill the SMs the dgrad chain
# leaves idle.  The side stream is joined back at the end of backward (autograd engine callback);
# gradient-ready callbacks fire inside the side-stream context, so the engines' events are
# recorded on the stream that actually produced the gradient.
_WGRAD_SIDE = os.environ.get("DTDS_WGRAD_STREAM", "1") != "0"
_side_streams: dict = {}
_join_queued: dict = {}


class _OnSideStream:
    def __init__(self, *tensors, params=()):
        self.tensors = [t for t in tensors if t is not None and t.is_cuda]
        # gradients written from several places (tied embeddings) stay on the main stream so the
        # overwrite-then-accumulate order of their producers is the program order
        shared = any(p is not None and getattr(p, "_grad_uses", 1) > 1 for p in params)
        # Only gradients written in place into a flat buffer (main_grad) may be produced here: the
        # engines / optimizers order their consumers after this stream.  A gradient handed back to
        # autograd as a tensor would be accumulated into .grad on the main stream with nothing
        # ordering that add after the side-stream GEMM.
        in_place = all(p is None or not p.requires_grad or getattr(p, "main_grad", None) is not None
                       for p in params)
        self.active = _WGRAD_SIDE and bool(self.tensors) and not shared and in_place

    def __enter__(self):
        if not self.active:
            return self
        dev = self.tensors[0].device
        side = _side_streams.get(dev)
        if side is None:
            side = _side_streams[dev] = torch.cuda.Stream(dev)
        cur = torch.cuda.current_stream(dev)
        side.wait_stream(cur)
        for t in self.tensors:
            t.record_stream(side)
        if not _join_queued.get(dev, False):
            _join_queued[dev] = True

            def join(dev=dev, side=side):
                torch.cuda.current_stream(dev).wait_stream(side)
                _join_queued[dev] = False
            try:
                torch.autograd.Variable._execution_engine.queue_callback(join)
            except RuntimeError:   # not inside a backward pass: join right away on exit
                _join_queued[dev] = False
                self._join_now = (dev, side)
        self._ctx = torch.cuda.stream(side)
        self._ctx.__enter__()
        return self

    def __exit__(self, *exc):
        if not self.active:
            return False
        self._ctx.__exit__(*exc)
        if getattr(self, "_join_now", None):
            dev, side = self._join_now
            torch.cuda.current_stream(dev).wait_stream(side)
        return False


def wgrad_side_stream(device):
    """The side stream weight gradients of `device` are produced on (None if never used, and None
    while a CUDA graph is being captured into which that stream has not been forked yet: it then
    holds no work of this step, and waiting on a stream outside the capture is an error)."""
    side = _side_streams.get(device)
    if side is not None and torch.cuda.is_current_stream_capturing():
        with torch.cuda.stream(side):
            if not torch.cuda.is_current_stream_capturing():
                return None
    return side


def _emit(param, compute):
    """compute(out, accumulate) -> tensor.  Returns the value autograd should see for `param`."""
    if param is None or not param.requires_grad:
        return None
    mg, acc = grad_target(param)
    if mg is not None:
        compute(mg, acc)
        grad_done(param)
        return None
    return compute(None, False)


def _wgrad(dy2, x2, param):
    """dW[N,K] = dy2[M,N]ᵀ · x2[M,K] emitted per the protocol above."""
    def compute(out, acc):
        if out is None:
            return K.gemm(dy2, x2, a_mn_major=True, b_mn_major=True, out_dtype=param.dtype)
        return K.gemm(dy2, x2, a_mn_major=True, b_mn_major=True, out=out, accumulate=acc)
    return _emit(param, compute)


def _bgrad(dy2, param):
    def compute(out, acc):
        if out is None:
            return K.colsum(dy2).to(param.dtype)
        return K.colsum(dy2, out=out, accumulate=acc)
    return _emit(param, compute)


# ------------------------------------------------------------------------------------ Linear
class LinearFn(torch.autograd.Function):
    """y = act(x·Wᵀ + b) (+ residual).  Forward: one GEMM with fused epilogue.  Backward: optional
    activation backward, dgrad GEMM (W consumed MN-major, no transpose), wgrad GEMM (both operands
    MN-major) written into the flat gradient buffer, bias column-sum."""

    @staticmethod
    def forward(ctx, x, weight, bias, act, residual, fork=False):
        shape = x.shape
        x2 = x.reshape(-1, shape[-1])
        N = weight.shape[0]
        ctx.fork = fork
        need_pre = act != ACT_NONE and any(ctx.needs_input_grad)
        pre = torch.empty(x2.shape[0], N, dtype=x2.dtype, device=x2.device) if need_pre else None
        res2 = residual.reshape(-1, N) if residual is not None else None
        if getattr(weight, "_ag_pending", False):
            # ZeRO-3: the weight has not been gathered; the GEMM kernel assembles it from the
            # owners' shards while it computes (all-gather fused with the consuming GEMM).
            assert pre is None, "fused all-gather GEMM does not save pre-activations"
            y = K.gemm_allgather(x2, weight.data, weight._ag_spec, bias=bias, act=act, residual=res2)
            weight._ag_pending = False
        else:
            y = K.gemm(x2, weight, bias=bias, act=act, residual=res2, pre_out=pre)
        ctx.save_for_backward(x2, weight, bias, pre)
        ctx.act = act
        ctx.has_res = residual is not None
        ctx.in_shape = shape
        if fork:   # second handle on the input for a skip connection (see `linear`)
            return y.view(*shape[:-1], N), x.view_as(x)
        return y.view(*shape[:-1], N)

    @staticmethod
    def backward(ctx, dy, dskip=None):
        x2, weight, bias, pre = ctx.saved_tensors
        N = weight.shape[0]
        dy2 = dy.reshape(-1, N)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dres = dy if ctx.has_res else None
        du = K.act_bwd(dy2, pre, ctx.act) if ctx.act != ACT_NONE else dy2
        dx = None
        if ctx.needs_input_grad[0]:
            # gradient of the skip branch: added in the epilogue of the dgrad GEMM
            dsk2 = _as2d(dskip, x2.shape[1]) if dskip is not None else None
            if getattr(weight, "_ag_pending", False):   # ZeRO-3 re-gather fused with the dgrad GEMM
                dx = K.gemm_allgather(du, weight.data, weight._ag_spec, b_mn_major=True).view(ctx.in_shape)
                weight._ag_pending = False
                if dsk2 is not None:
                    dx = dx + dskip
            else:
                dx = K.gemm(du, weight, b_mn_major=True, residual=dsk2).view(ctx.in_shape)
        with _OnSideStream(du, x2, params=(weight, bias)):
            dw = _wgrad(du, x2, weight) if ctx.needs_input_grad[1] else None
            db = _bgrad(du, bias) if (bias is not None and ctx.needs_input_grad[2]) else None
        return dx, dw, db, None, dres, None


FUSE_SKIP_GRADS = os.environ.get("DTDS_FUSE_SKIP_GRADS", "1") != "0"


def _as2d(t, cols):
    t2 = t.reshape(-1, cols)
    return t2 if t2.is_contiguous() else t2.contiguous()


def linear(x, weight, bias=None, act=ACT_NONE, residual=None, fork=False):
    """``fork=True`` returns ``(y, x_skip)``: x_skip is x again, to be used as the residual of a
    later op.  Its gradient comes back to this function's backward as a separate argument and is
    added in the epilogue of the dgrad GEMM, instead of autograd summing the two branches of x
    with a standalone elementwise kernel.  (FUSE_SKIP_GRADS = False: plain autograd, for A/B.)"""
    if fork and not FUSE_SKIP_GRADS:
        return LinearFn.apply(x, weight, bias, act, residual, False), x
    return LinearFn.apply(x, weight, bias, act, residual, fork)


# --------------------------------------------------------------------------------------- MLP
class MLPFn(torch.autograd.Function):
    """f = act(x·W1ᵀ + b1)·W2ᵀ + b2 (+ residual).  The backward of the activation is fused into
    the epilogue of the W2 dgrad GEMM (du = (df·W2) ⊙ act'(u)); no standalone elementwise pass."""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2, act, residual, fork=False):
        shape = x.shape
        x2 = x.reshape(-1, shape[-1])
        M, F = x2.shape[0], w1.shape[0]
        u = torch.empty(M, F, dtype=x2.dtype, device=x2.device)
        a = K.gemm(x2, w1, bias=b1, act=act, pre_out=u)
        res2 = residual.reshape(-1, w2.shape[0]) if residual is not None else None
        f = K.gemm(a, w2, bias=b2, residual=res2)
        ctx.save_for_backward(x2, w1, b1, w2, b2, u, a)
        ctx.act = act
        ctx.has_res = residual is not None
        ctx.in_shape = shape
        if fork:
            return f.view(*shape[:-1], w2.shape[0]), x.view_as(x)
        return f.view(*shape[:-1], w2.shape[0])

    @staticmethod
    def backward(ctx, df, dskip=None):
        x2, w1, b1, w2, b2, u, a = ctx.saved_tensors
        df2 = df.reshape(-1, w2.shape[0])
        if not df2.is_contiguous():
            df2 = df2.contiguous()
        dres = df if ctx.has_res else None
        du = K.gemm(df2, w2, b_mn_major=True, act=ctx.act, dact_in=u)
        # after the dgrad that reads w2: a gradient-ready callback may update w2 right away
        with _OnSideStream(df2, a, params=(w2, b2)):
            dw2 = _wgrad(df2, a, w2)
            db2 = _bgrad(df2, b2) if b2 is not None else None
        dx = None
        if ctx.needs_input_grad[0]:
            dsk2 = _as2d(dskip, x2.shape[1]) if dskip is not None else None
            dx = K.gemm(du, w1, b_mn_major=True, residual=dsk2).view(ctx.in_shape)
        with _OnSideStream(du, x2, params=(w1, b1)):
            dw1 = _wgrad(du, x2, w1)
            db1 = _bgrad(du, b1) if b1 is not None else None
        return dx, dw1, db1, dw2, db2, None, dres, None


def mlp(x, w1, b1, w2, b2, act, residual=None, fork=False):
    """``fork=True``: see `linear`."""
    if fork and not FUSE_SKIP_GRADS:
        return MLPFn.apply(x, w1, b1, w2, b2, act, residual, False), x
    return MLPFn.apply(x, w1, b1, w2, b2, act, residual, fork)


# --------------------------------------------------------------------------------- LayerNorm
class LayerNormFn(torch.autograd.Function):
    """y = LN(dropout(x) + residual).  One kernel forward (dropout mask recomputed from its seed
    in backward, never stored), one kernel + a tiny finalize backward."""

    @staticmethod
    def forward(ctx, x, residual, gamma, beta, eps, p_drop, training, fork=False):
        p = p_drop if training else 0.0
        if fork:
            assert residual is None and p == 0.0, "fork: plain LayerNorm(x) only"
        seed = next_seed() if p > 0.0 else 0
        y, z, mean, rstd = K.layernorm_fwd(x, residual, gamma, beta, eps, p, seed)
        ctx.save_for_backward(z, mean, rstd, gamma, beta)
        ctx.p, ctx.seed = p, seed
        ctx.has_res = residual is not None
        if fork:   # pre-LN blocks: x also feeds the skip connection (see `linear`)
            return y, x.view_as(x)
        return y

    @staticmethod
    def backward(ctx, dy, dskip=None):
        z, mean, rstd, gamma, beta = ctx.saved_tensors
        if not dy.is_contiguous():
            dy = dy.contiguous()
        mg_g, acc_g = grad_target(gamma)
        mg_b, acc_b = grad_target(beta)
        direct = mg_g is not None and mg_b is not None and acc_g == acc_b
        if direct and _WGRAD_SIDE and K.layernorm_bwd_splittable(z):
            # dz feeds the next backward op; dgamma/dbeta feed nobody until the optimizer: they
            # run on the weight-gradient side stream like the wgrad GEMMs.
            dz, dxd, _, _ = K.layernorm_bwd(dy, z, mean, rstd, gamma, dz_extra=dskip, p_drop=ctx.p,
                                            seed=ctx.seed, want_dx_drop=True, parts="dx")
            with _OnSideStream(dy, z, mean, rstd, params=(gamma, beta)):
                K.layernorm_bwd(dy, z, mean, rstd, gamma, dgamma_out=mg_g, dbeta_out=mg_b,
                                accumulate=acc_g, parts="params")
                grad_done(gamma)
                grad_done(beta)
            return dxd, (dz if ctx.has_res else None), None, None, None, None, None, None
        dz, dxd, dg, db = K.layernorm_bwd(
            dy, z, mean, rstd, gamma, dz_extra=dskip, p_drop=ctx.p, seed=ctx.seed,
            dgamma_out=mg_g if direct else None, dbeta_out=mg_b if direct else None,
            accumulate=acc_g if direct else False, want_dx_drop=True)
        if direct:
            grad_done(gamma)
            grad_done(beta)
            dg = db = None
        else:
            dg = _emit(gamma, lambda out, acc: dg if out is None else (out.add_(dg) if acc else out.copy_(dg)))
            db = _emit(beta, lambda out, acc: db if out is None else (out.add_(db) if acc else out.copy_(db)))
        return dxd, (dz if ctx.has_res else None), dg, db, None, None, None, None


def layer_norm(x, gamma, beta, eps=1e-5, residual=None, p_drop=0.0, training=False, fork=False):
    """``fork=True`` (plain LN only): returns ``(y, x_skip)``; the skip gradient is added inside
    the backward kernel (dz_extra) rather than by a standalone add."""
    if fork and not FUSE_SKIP_GRADS:
        return LayerNormFn.apply(x, residual, gamma, beta, eps, p_drop, training, False), x
    return LayerNormFn.apply(x, residual, gamma, beta, eps, p_drop, training, fork)


# --------------------------------------------------------------------------------- attention
class AttentionFn(torch.autograd.Function):
    """ctx[B*S, H] = softmax(scale·QKᵀ + bias)·V from the fused QKV activations [B*S, 3H]."""

    @staticmethod
    def forward(ctx, qkv, B, S, heads, scale, causal, key_bias, alibi, p_drop, training):
        p = p_drop if training else 0.0
        seed = next_seed() if p > 0.0 else 0
        out, lse = K.attention_fwd(qkv, B, S, heads, scale=scale, causal=causal, key_bias=key_bias,
                                   alibi=alibi, p_drop=p, seed=seed)
        ctx.save_for_backward(qkv, out, lse, key_bias, alibi)
        ctx.cfg = (B, S, heads, scale, causal, p, seed)
        return out

    @staticmethod
    def backward(ctx, dout):
        qkv, out, lse, key_bias, alibi = ctx.saved_tensors
        B, S, heads, scale, causal, p, seed = ctx.cfg
        dqkv = K.attention_bwd(qkv, out, dout, lse, B, S, heads, scale=scale, causal=causal,
                               key_bias=key_bias, alibi=alibi, p_drop=p, seed=seed)
        return (dqkv,) + (None,) * 9


def attention(qkv, B, S, heads, *, scale=None, causal=False, key_bias=None, alibi=None,
              p_drop=0.0, training=False):
    H = qkv.shape[-1] // 3
    if scale is None:
        scale = 1.0 / math.sqrt(H // heads)
    if qkv.is_cuda and S % 128 != 0:
        # the tcgen05 kernels work on 128-row blocks: other lengths run padded, with the extra keys
        # masked out and the extra query rows dropped
        return attention_padded(qkv, B, S, heads, scale=scale, causal=causal, key_bias=key_bias, alibi=alibi,
                                p_drop=p_drop, training=training)
    return AttentionFn.apply(qkv, B, S, heads, scale, causal, key_bias, alibi, p_drop, training)


def attention_padded(qkv, B, S, heads, *, scale, causal=False, key_bias=None, alibi=None, p_drop=0.0,
                     training=False, multiple: int = 128):
    """Attention for a sequence length that is not a multiple of the kernels' 128-row block: pads
    Q/K/V rows with zeros up to the next multiple, gives the padded keys a -1e9 bias (exp
    underflows to exactly 0) and returns the first S rows.  Padding and slicing are ordinary
    autograd ops around the fused function, so the backward needs nothing extra."""
    Sp = (S + multiple - 1) // multiple * multiple
    H3 = qkv.shape[-1]
    padded = torch.nn.functional.pad(qkv.reshape(B, S, H3), (0, 0, 0, Sp - S)).reshape(B * Sp, H3)
    kb = torch.zeros(B, Sp, dtype=torch.float32, device=qkv.device)
    kb[:, S:] = -1e9
    if key_bias is not None:
        kb[:, :S] = key_bias.to(torch.float32)
    out = AttentionFn.apply(padded, B, Sp, heads, scale, causal, kb, alibi, p_drop, training)
    return out.reshape(B, Sp, H3 // 3)[:, :S].reshape(B * S, H3 // 3)


# --------------------------------------------------------------------------------- embedding
class EmbeddingFn(torch.autograd.Function):
    """word[ids] + pos[0..S) + type[0]: one gather kernel forward; backward = atomic scatter-add
    into the word table gradient, batch reduction for positions, column sum for the type row."""

    @staticmethod
    def forward(ctx, ids, word, pos, typ, pos_offset):
        out = K.embedding_fwd(ids, word, pos, typ, None, pos_offset)
        ctx.save_for_backward(ids)
        ctx.params = (word, pos, typ)
        ctx.pos_offset = pos_offset
        return out

    @staticmethod
    def backward(ctx, dout):
        (ids,) = ctx.saved_tensors
        word, pos, typ = ctx.params
        B, S = ids.shape
        H = dout.shape[-1]
        if not dout.is_contiguous():
            dout = dout.contiguous()

        def word_compute(out, acc):
            tgt = out
            if tgt is None:
                tgt = torch.zeros_like(word)
            elif not acc:
                tgt.zero_()
            K.embedding_bwd(ids, dout, tgt, None)
            return tgt

        def pos_compute(out, acc):
            tgt = out
            if tgt is None:
                tgt = torch.zeros_like(pos)
            elif not acc:
                tgt.zero_()
            K.embedding_bwd(ids, dout, None, tgt, pos_offset=ctx.pos_offset, accumulate_pos=True)
            return tgt

        def type_compute(out, acc):
            tgt = out
            if tgt is None:
                tgt = torch.zeros_like(typ)
            elif not acc:
                tgt.zero_()
            K.colsum(dout.view(B * S, H), out=tgt[0], accumulate=True)
            return tgt

        dword = _emit(word, word_compute)
        dpos = _emit(pos, pos_compute) if pos is not None else None
        dtyp = _emit(typ, type_compute) if typ is not None else None
        return None, dword, dpos, dtyp, None


def embedding(ids, word, pos=None, typ=None, pos_offset=0):
    return EmbeddingFn.apply(ids, word, pos, typ, pos_offset)


# ------------------------------------------------------------------------------------- losses
class CrossEntropyFn(torch.autograd.Function):
    """Mean cross-entropy with ignore_index over logits [R, V] (row stride may be padded).

    Forward and backward are one pass over the logits: the gradient is written in place into the
    logits buffer, so the op needs no second [R, V] tensor.  The upstream gradient is assumed to
    be the scalar ``gscale`` given at call time (1.0 for a plain ``loss.backward()``)."""

    @staticmethod
    def forward(ctx, logits2, labels, vocab, ignore_index, shift, seq, gscale):
        want_grad = ctx.needs_input_grad[0]
        loss, d = K.cross_entropy(logits2, labels, vocab=vocab, ignore_index=ignore_index,
                                  shift=shift, seq=seq, want_grad=want_grad, inplace_grad=True,
                                  gscale=gscale)
        ctx.d = d
        return loss.clone()

    @staticmethod
    def backward(ctx, grad_out):
        d, ctx.d = ctx.d, None
        return d, None, None, None, None, None, None


def cross_entropy(logits, labels, *, ignore_index=-100, shift=False, gscale=1.0):
    """logits [..., V] (last dim contiguous; rows may be padded), labels [...] int64."""
    V = logits.shape[-1]
    if logits.dim() == 3:
        seq = logits.shape[1]
        l2 = logits.reshape(-1, V) if logits.is_contiguous() else \
            logits.as_strided((logits.shape[0] * seq, V), (logits.stride(1), 1), logits.storage_offset())
    else:
        seq = logits.shape[0]
        l2 = logits
    if l2.is_cuda and (l2.stride(0) % 8 != 0):
        pad = (V + 7) // 8 * 8
        buf = torch.zeros(l2.shape[0], pad, dtype=l2.dtype, device=l2.device)
        buf[:, :V] = l2
        l2 = buf[:, :V]
    return CrossEntropyFn.apply(l2, labels.reshape(-1), V, ignore_index, shift, seq, gscale)


class LMHeadLossFn(torch.autograd.Function):
    """loss = CE(x·Wᵀ + b, labels): vocabulary projection + cross-entropy as one op.  The logits
    live in one padded [M, ld] buffer that the CE kernel turns into their own gradient in place;
    backward is then the dgrad / wgrad / bias column-sum of the projection."""

    @staticmethod
    def forward(ctx, x, weight, bias, labels, ignore_index, shift, seq, gscale, n_valid=None):
        x2 = x.reshape(-1, x.shape[-1])
        M, V = x2.shape[0], weight.shape[0]
        ld = (V + 63) // 64 * 64
        buf = torch.empty(M, ld, dtype=x2.dtype, device=x2.device)
        if ld != V:
            buf[:, V:].zero_()
        logits = buf[:, :V]
        K.gemm(x2, weight, bias=bias, out=logits)
        want_grad = any(ctx.needs_input_grad)
        loss, d = K.cross_entropy(buf, labels.reshape(-1), vocab=V, ignore_index=ignore_index,
                                  shift=shift, seq=seq, want_grad=want_grad, inplace_grad=True,
                                  gscale=gscale, n_valid=n_valid)
        ctx.save_for_backward(x2, weight, bias)
        ctx.d = d[:, :V] if d is not None else None
        ctx.in_shape = x.shape
        return loss.clone()

    @staticmethod
    def backward(ctx, grad_out):
        x2, weight, bias = ctx.saved_tensors
        d, ctx.d = ctx.d, None
        dx = K.gemm(d, weight, b_mn_major=True).view(ctx.in_shape) if ctx.needs_input_grad[0] else None
        with _OnSideStream(d, x2, params=(weight, bias)):
            dw = _wgrad(d, x2, weight)
            db = _bgrad(d, bias) if bias is not None else None
        return dx, dw, db, None, None, None, None, None, None


def lm_head_loss(x, weight, bias, labels, *, ignore_index=-100, shift=False, seq=None, gscale=1.0, n_valid=None):
    """`n_valid`: 1-element float device tensor to normalise by (default: the rows' own count)."""
    seq = seq or (x.shape[1] if x.dim() == 3 else x.shape[0])
    return LMHeadLossFn.apply(x, weight, bias, labels, ignore_index, shift, seq, gscale, n_valid)


class MSEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, target):
        loss, dx = K.mse(x, target, want_grad=x.requires_grad)
        ctx.dx = dx
        return loss

    @staticmethod
    def backward(ctx, grad_out):
        dx, ctx.dx = ctx.dx, None
        return dx, None


def mse_loss(x, target):
    return MSEFn.apply(x, target)
