"""ZeRO stage 3: parameter sharding with just-in-time materialisation.

What DeepSpeed's ``partition_parameters`` / ``partitioned_param_coordinator`` do for the
reference (zero_dp_training.py ``--stage 3``), re-designed for NVSwitch:

  * between uses a rank keeps only its 1/n slice of every bucket ("shard"), in a symmetric buffer
    that every peer maps;
  * parameters are materialised per *block* (each element of a ``ModuleList`` — one transformer
    layer — plus the remaining children of the root): a pre-forward hook pulls the block's slices
    straight out of the owners' shard buffers over NVLink (one ``copy_segments`` launch, peer
    loads, no barrier, no staging all-gather buffer on the owner side), a post-forward hook drops
    them; the same happens around the block's backward.  The next block is prefetched on a side
    stream while the current one computes;
  * parameters used by several blocks (tied embeddings) stay materialised for the whole step;
  * after a block's backward its bucket runs the fused reduce-scatter + Adam kernel, which writes
    the updated bf16 slice into the owner's shard (no all-gather at all in stage 3).

On CPU/gloo or ``DTDS_COMM=nccl`` materialisation uses ``all_gather_into_tensor`` on the shards.
"""
from __future__ import annotations

import ctypes

import torch
from torch import nn

from .. import comm
from ..ops import _lib


class _Block:
    __slots__ = ("name", "module", "params", "buckets", "live", "event", "lead")

    def __init__(self, name, module, params):
        self.name, self.module, self.params = name, module, params
        self.buckets, self.live, self.event = [], False, None
        self.lead = None  # parameter whose gather is fused into its consuming GEMM


class Zero3ParamManager:
    def __init__(self, engine):
        self.e = engine
        f = engine.flat
        self.flat = f
        self.dev = engine.device
        self.world, self.rank = engine.world, engine.rank
        total_own = sum(b.per for b in engine.buckets)
        esz = f.param_flat.element_size()
        self.custom = engine._custom and self.world > 1
        if self.custom:
            ctx = comm.symm_context(engine.group)
            self.shard_buf = ctx.new_buffer(total_own * esz)
            self.shards = self.shard_buf.tensor(f.param_dtype, total_own)
        else:
            self.shard_buf = None
            self.shards = torch.empty(total_own, dtype=f.param_dtype, device=self.dev)
        self.refresh_shards_from_full()
        self.grad_buf = self.grads = None   # set by _build_gradient_arena once the blocks are known
        # ---- blocks
        root = engine.module
        self.blocks: list[_Block] = []
        claimed = set()
        pinned = [p for p in f.params if getattr(p, "_grad_uses", 1) > 1]
        claimed.update(id(p) for p in pinned)

        def add(name, mod):
            ps = [p for p in mod.parameters() if id(p) not in claimed and p.requires_grad]
            if ps:
                claimed.update(id(p) for p in ps)
                self.blocks.append(_Block(name, mod, ps))

        for cname, child in root.named_children():
            if isinstance(child, nn.ModuleList):
                for i, sub in enumerate(child):
                    add(f"{cname}.{i}", sub)
            else:
                add(cname, child)
        rest = [p for p in f.params if id(p) not in claimed]
        self.pinned = _Block("<root>", root, pinned + rest)
        bucket_of = engine._bucket_of
        for blk in self.blocks + [self.pinned]:
            seen = {}
            for p in blk.params:
                b = bucket_of[id(p)]
                seen[b.index] = b
            blk.buckets = list(seen.values())
        self._build_gradient_arena()
        # all-gather fused with the first consuming GEMM: one flagged Linear weight per block
        self.fuse_gather = self.custom and bool(engine.config.get("zero_optimization", {}).get(
            "fuse_allgather_gemm", True))
        if self.fuse_gather:
            for blk in self.blocks:
                for p in blk.params:
                    if getattr(p, "_ag_ok", False) and p.dim() == 2 and (p._orig_shape[1] * 2) % 16 == 0:
                        blk.lead = p
                        break
        # ---- release the full replica: from here on parameters point at per-block gather buffers
        self._full_released = False
        self._side = torch.cuda.Stream(self.dev) if self.dev.type == "cuda" else None
        self._order: list[_Block] = []
        self._placeholder = torch.empty(0, dtype=f.param_dtype, device=self.dev)
        self._release_full_replica()
        for i, blk in enumerate(self.blocks):
            blk.module.register_forward_pre_hook(self._pre(blk, i, +1))
            blk.module.register_forward_hook(self._post(blk))
            blk.module.register_full_backward_pre_hook(self._pre_bwd(blk, i))
            blk.module.register_full_backward_hook(self._post_bwd(blk))
        root.register_forward_pre_hook(lambda m, a: self._materialize(self.pinned))
        if self.dev.type == "cuda":
            torch.cuda.synchronize(self.dev)
        if self.world > 1:
            comm.barrier(engine.group)  # every rank's shards are in place before anyone pulls

    # ------------------------------------------------------------------------------------
    def _build_gradient_arena(self) -> None:
        """Stage 3 shards the gradients too (DeepSpeed stage 3, zero_dp_training.py:41 `--stage 3`):
        a bucket's gradients are dead once its reduce-scatter + Adam kernel has run, which happens
        right behind the backward of the block that produced them.  So instead of a model-sized
        buffer the gradients live in an arena of `grad_ring_blocks` slots, each large enough for the
        buckets of one block, that the blocks take turns using in backward order, plus fixed space
        for the buckets of parameters that collect gradients over the whole backward (tied
        embeddings, root-level parameters).  A bucket belongs to the block whose backward writes
        into it first; before that block's backward starts, the stream waits for the kernels of the
        buckets that used its slot `grad_ring_blocks` blocks earlier (events of the same step)."""
        e, f = self.e, self.flat
        zcfg = e.config.get("zero_optimization", {}) or {}
        depth = max(2, int(zcfg.get("grad_ring_blocks", 3)))
        block_of = {}
        for bi, blk in enumerate(self.blocks):
            for p in blk.params:
                block_of[id(p)] = bi
        fixed, groups = [], {}
        for b in e.buckets:   # b.params[0] is the bucket's last parameter = its first in backward
            owners = [block_of.get(id(p)) for p in b.params]
            if any(o is None for o in owners):
                fixed.append(b)
            else:
                groups.setdefault(max(owners), []).append(b)   # latest block in forward order runs first
        order = sorted(groups, reverse=True)                   # blocks in backward order
        slot = max((sum(b.hi - b.lo for b in groups[k]) for k in order), default=0)
        slots = min(depth, len(order))
        off = 0
        for b in fixed:
            b.goff, b.prev_user = off, None
            off += b.hi - b.lo
        base = off
        self._slot_waits = {}   # block index -> buckets whose kernels must have finished before its backward
        for n, k in enumerate(order):
            o = base + (n % slots) * slot
            for b in groups[k]:
                b.goff = o
                o += b.hi - b.lo
            prev = groups[order[n - slots]] if n >= slots else []
            for b in groups[k]:
                b.prev_user = prev[0] if prev else None
            self._slot_waits[k] = list(prev)
        total = base + slots * slot
        esz = torch.empty((), dtype=f.grad_dtype).element_size()
        if self.custom and f.grad_dtype == torch.bfloat16:
            buf = f._try_multicast(total * esz, e.group) if getattr(e, "_want_mc", False) else None
            self.grad_buf = buf if buf is not None else comm.symm_context(e.group).new_buffer(total * esz)
            self.grads = self.grad_buf.tensor(f.grad_dtype, total)
        else:
            self.grads = torch.zeros(total, dtype=f.grad_dtype, device=self.dev)
        for b in e.buckets:
            for p in b.params:
                o = b.goff + (p._flat_offset - b.lo)
                p.main_grad = self.grads[o:o + p._numel].view(p._orig_shape)
        f.grad_flat = f.grad_buf = None

    def grad_view(self, b) -> torch.Tensor:
        return self.grads[b.goff:b.goff + (b.hi - b.lo)]

    def gradient_bytes(self) -> int:
        return self.grads.numel() * self.grads.element_size()

    def _wait_slots(self, blk_index: int) -> None:
        """Before a block's backward writes gradients: the previous users of its arena slot must have
        been drained (stream-ordered; the weight-gradient side stream forks from the current stream
        after this point).  Their kernels were enqueued earlier in this same backward, so under graph
        capture these are waits on events of the capture itself."""
        if self.dev.type != "cuda":
            return
        cur = torch.cuda.current_stream(self.dev)
        for b in self._slot_waits.get(blk_index, ()):
            if b.done_event is not None and b.launched:
                cur.wait_event(b.done_event)

    def shard_view(self, b) -> torch.Tensor:
        return self.shards[b.state_off:b.state_off + b.per]

    def refresh_shards_from_full(self) -> None:
        f = self.flat
        for b in self.e.buckets:
            s = b.lo + self.rank * b.per
            self.shard_view(b).copy_(f.param_flat[s:s + b.per])

    def _release_full_replica(self) -> None:
        """Drops the full parameter buffer; every parameter becomes a zero-size placeholder until
        its block is materialised.  This is where the 1/n parameter memory comes from."""
        f = self.flat
        for p in f.params:
            p.data = self._placeholder
        f.param_flat = None
        f.param_buf = None
        self._full_released = True
        if self.dev.type == "cuda":
            torch.cuda.empty_cache()

    def resident_parameter_bytes(self) -> int:
        """Bytes of parameter storage this rank holds between steps (its shards only)."""
        return self.shards.numel() * self.shards.element_size()

    # ------------------------------------------------------------------------------------
    def _segments(self, b, lo: int, hi: int):
        """(source address, byte offset inside bucket b, nbytes) pieces covering flat elements
        [lo, hi) of bucket b, in memory order, one piece per owning rank."""
        esz = self.shards.element_size()
        out = []
        for r in range(self.world):
            s0, s1 = b.lo + r * b.per, b.lo + (r + 1) * b.per
            a, z = max(lo, s0), min(hi, s1)
            if a < z:
                src = self.shard_buf.ptrs[r] + (b.state_off + (a - s0)) * esz
                out.append((src, (a - b.lo) * esz, (z - a) * esz))
        return out

    def _gather_bucket(self, b, skip=None) -> torch.Tensor:
        """Full copy of bucket `b` assembled from all ranks' shards (transient).  `skip` = flat
        element range [lo, hi) left for a fused all-gather GEMM to fill."""
        f = self.flat
        n = b.hi - b.lo
        full = torch.empty(n, dtype=self.shards.dtype, device=self.dev)
        esz = full.element_size()
        if self.world == 1:
            full.copy_(self.shard_view(b))
        elif self.custom:
            L = _lib.lib()
            if skip is None or skip[1] <= b.lo or skip[0] >= b.hi:
                segs = self._segments(b, b.lo, b.hi)
            else:
                segs = self._segments(b, b.lo, skip[0]) + self._segments(b, skip[1], b.hi)
            ns = len(segs)
            srcs = (ctypes.c_ulonglong * ns)(*[s for s, _, _ in segs])
            dsts = (ctypes.c_ulonglong * ns)(*[full.data_ptr() + o for _, o, _ in segs])
            sizes = (ctypes.c_ulonglong * ns)(*[nb for _, _, nb in segs])
            with comm.comms_logger.record("all_gather(pull)", sum(nb for _, _, nb in segs), self.world, full):
                _lib.check(L.dtds_copy_segments(srcs, dsts, sizes, ns, 0, _lib.stream_ptr()),
                           "dtds_copy_segments")
        else:
            comm.all_gather_into_tensor(full, self.shard_view(b).contiguous(), group=self.e.group)
        return full

    def _materialize(self, blk: _Block, stream=None) -> None:
        if blk.live:
            if blk.event is not None:  # prefetched on the side stream: order it before our use
                torch.cuda.current_stream(self.dev).wait_event(blk.event)
                blk.event = None
            return
        ctx = torch.cuda.stream(stream) if stream is not None else _Null()
        lead = blk.lead
        skip = (lead._flat_offset, lead._flat_offset + lead._numel) if lead is not None else None
        with ctx:
            fulls = {b.index: self._gather_bucket(b, skip) for b in blk.buckets}
        for p in blk.params:
            b = self.e._bucket_of[id(p)]
            o = p._flat_offset - b.lo
            p.data = fulls[b.index][o:o + p._numel].view(p._orig_shape)
        if lead is not None:
            # the consuming GEMM pulls this weight itself (see ops/functional.py: LinearFn)
            if getattr(lead, "_ag_spec", None) is None:
                from ..ops.kernels import GatherSpec
                b = self.e._bucket_of[id(lead)]
                esz = self.shards.element_size()
                pieces = self._segments(b, skip[0], skip[1])
                segs = [(s, nb) for s, _, nb in pieces]
                row_bytes = lead._orig_shape[1] * esz
                # where this rank's own slice starts inside the weight (0 if it owns none of it)
                mine = self.shard_buf.ptrs[self.rank]
                span = self.shards.numel() * esz
                local_begin = next((o - (skip[0] - b.lo) * esz for s, o, nb in pieces if mine <= s < mine + span), 0)
                lead._ag_spec = GatherSpec(segs, lead._numel * esz, self.dev, chunk_rows=32, row_bytes=row_bytes,
                                           local_begin=max(0, local_begin))
            lead._ag_pending = True
        if stream is not None:
            blk.event = torch.cuda.Event()
            blk.event.record(stream)
            for t in fulls.values():  # buffers allocated on the side stream, used on the main one
                t.record_stream(torch.cuda.current_stream(self.dev))
        blk.live = True

    def _release(self, blk: _Block) -> None:
        if not blk.live or blk is self.pinned:
            return
        for p in blk.params:
            p.data = self._placeholder
        if blk.lead is not None:
            blk.lead._ag_pending = False
        blk.live = False
        blk.event = None

    def _prefetch(self, idx: int) -> None:
        if self._side is None or not self.e.overlap_comm or not (0 <= idx < len(self.blocks)):
            return
        blk = self.blocks[idx]
        if blk.live:
            return
        self._side.wait_stream(torch.cuda.current_stream(self.dev))
        self._materialize(blk, stream=self._side)

    # hooks ------------------------------------------------------------------------------
    def _pre(self, blk, idx, direction):
        def hook(module, args):
            self._materialize(blk)
            self._prefetch(idx + direction)
        return hook

    def _post(self, blk):
        def hook(module, args, output):
            self._release(blk)
        return hook

    def _pre_bwd(self, blk, idx):
        def hook(module, grad_output):
            self._wait_slots(idx)
            self._materialize(blk)
            self._prefetch(idx - 1)
        return hook

    def _post_bwd(self, blk):
        def hook(module, grad_input, grad_output):
            # If no input of the block requires grad (the embedding block) PyTorch fires this hook
            # BEFORE the block's backward has run; releasing then would pull the parameters out
            # from under it.  Such blocks are dropped at the end of the step instead.
            if any(g is not None for g in grad_input):
                self._release(blk)
        return hook

    def release_all(self) -> None:
        for blk in self.blocks:
            self._release(blk)
        for p in self.pinned.params:
            p.data = self._placeholder
        self.pinned.live = False

    def gather_all(self) -> None:
        """Materialises every parameter (checkpoint export / evaluation outside the engine)."""
        for blk in self.blocks + [self.pinned]:
            self._materialize(blk)
            if blk.lead is not None and getattr(blk.lead, "_ag_pending", False):
                self._complete_lead(blk.lead)

    def _complete_lead(self, p) -> None:
        """Plain gather of a weight that was left for a fused all-gather GEMM."""
        b = self.e._bucket_of[id(p)]
        segs = self._segments(b, p._flat_offset, p._flat_offset + p._numel)
        ns = len(segs)
        base = p.data.data_ptr() - (p._flat_offset - b.lo) * self.shards.element_size()
        srcs = (ctypes.c_ulonglong * ns)(*[s for s, _, _ in segs])
        dsts = (ctypes.c_ulonglong * ns)(*[base + o for _, o, _ in segs])
        sizes = (ctypes.c_ulonglong * ns)(*[nb for _, _, nb in segs])
        _lib.check(_lib.lib().dtds_copy_segments(srcs, dsts, sizes, ns, 0, _lib.stream_ptr()),
                   "dtds_copy_segments")
        p._ag_pending = False


class _Null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
