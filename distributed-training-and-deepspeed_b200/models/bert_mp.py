"""``BertModelWithMP``: BERT split into contiguous module groups across devices (naive model
parallelism) with a GPipe adapter — the first-party engine of the reference
(model/bert_mp.py:14-130), rebuilt on the fused kernels and NCCL p2p.

Same construction as the reference: ``[embeddings] + encoders + [head]`` (14 modules for
bert-base) divided with ``np.array_split`` semantics, an *untied* decoder, only ``input_ids``
fed (positions 0..S-1, token type 0, no attention mask), ``state_dict`` prefixes ``embeddings``,
``encoders.N`` and ``head``.  Same surface: ``embedding_device``, ``encoder_devices``,
``head_device``, ``device_idle_time``, ``to_pipeline(chunks)``, ``forward(input_ids)``.

Two execution modes:
  * one process per stage (``torch.distributed`` initialised with world_size == device_count):
    each rank builds only its own modules; ``forward_backward`` runs the GPipe schedule with NCCL
    send/recv overlapped with micro-batch compute (parallel/pipeline.py).  This is the mode the
    CLI uses on GPUs.
  * one process, several devices (or CPU): modules are placed on ``cuda:i`` and the activation
    hops with ``.to(device)`` exactly like the reference; ``to_pipeline`` then runs micro-batches
    back to back through the same placement.
"""
from __future__ import annotations

from datetime import datetime

import torch
import torch.distributed as dist
from torch import nn

from ..parallel.pipeline import IdleTracker, PipelineStage, split_ranges
from .bert import BertEmbeddings, BertLayer, BertOnlyMLMHead
from .config import ModelConfig


def _device_type() -> str:
    return "cuda" if torch.cuda.is_available() else "cpu"


def _device_count() -> int:
    return torch.cuda.device_count() if torch.cuda.is_available() else 1


class _ValueScale(torch.autograd.Function):
    """Scales the reported value of a loss whose gradient scale is already baked in."""

    @staticmethod
    def forward(ctx, x, w):
        return x * w

    @staticmethod
    def backward(ctx, g):
        return g, None


class LocalValue:
    """Return type of the pipelined model: mirrors the ``RRef.local_value()`` call the reference
    script makes on ``Pipe`` outputs (model_parallel_training.py:69-70)."""

    def __init__(self, value):
        self._value = value

    def local_value(self):
        return self._value


class BertModelWithMP(nn.Module):
    def __init__(self, config: ModelConfig, device_count: int | None = None, verbose: bool = False,
                 dtype: torch.dtype | None = None, process_group=None):
        super().__init__()
        self.config = config.replace(tie_word_embeddings=False)
        self.verbose = verbose
        self.group = process_group
        self.distributed = dist.is_available() and dist.is_initialized() and dist.get_world_size(process_group) > 1
        device_type = _device_type()
        n_modules = self.config.num_hidden_layers + 2
        if self.distributed:
            device_count = dist.get_world_size(process_group)
            self.stage = dist.get_rank(process_group)
        else:
            device_count = device_count or _device_count()
            self.stage = 0
            if dist.is_available() and dist.is_initialized():
                device_count = 1
        if not self.distributed or self.stage == 0:
            print("Using device: {}".format(device_type))
        available = _device_count() if not self.distributed else device_count
        if device_count <= 1:
            if not self.distributed:
                print(f"Only one {device_type} device is in use. Training will be done without Model Parallelism.")
        elif device_count > available:
            print(f"Cannot use {device_count} {device_type} devices. Only {available} devices are available. "
                  f"Training will be done using {available} devices.")
            device_count = available
        device_count = max(1, min(device_count, n_modules))
        self.device_count = device_count
        self.dtype = dtype or (torch.bfloat16 if device_type == "cuda" else torch.float32)
        groups = split_ranges(n_modules, device_count)
        self.group_ranges = groups
        # device of every module index (embedding = 0, encoders = 1..L, head = L+1)
        self.devices = []
        for gi, rng in enumerate(groups):
            dev = torch.device("cuda", gi) if device_type == "cuda" else torch.device("cpu")
            self.devices.extend([dev] * len(rng))
        mine = set(groups[self.stage]) if self.distributed else set(range(n_modules))
        self.local_device = (torch.device("cuda", torch.cuda.current_device()) if device_type == "cuda"
                             else torch.device("cpu"))

        def place(idx, make):
            if idx not in mine:
                return nn.Identity()
            dev = self.local_device if self.distributed else self.devices[idx]
            return make().to(device=dev, dtype=self.dtype)

        L = self.config.num_hidden_layers
        self.embeddings = place(0, lambda: BertEmbeddings(self.config))
        self.encoders = nn.ModuleList([place(1 + i, lambda: BertLayer(self.config)) for i in range(L)])
        self.head = place(L + 1, lambda: BertOnlyMLMHead(self.config))
        self._mine = sorted(mine)
        # idle tracking: one tracker per device this process drives
        self.trackers = {}
        if self.distributed:
            self.trackers[self.stage] = IdleTracker(self.local_device, verbose, self.log)
        else:
            for gi in range(device_count):
                dev = self.devices[groups[gi][0]]
                self.trackers[gi] = IdleTracker(dev, verbose, self.log)
        self.device_idle_time = {d: (0, 0) for d in range(device_count)}
        self._chunks = 1
        self._pipe = None

    # ------------------------------------------------------------------ reference surface
    @property
    def embedding_device(self):
        return self.devices[0]

    @property
    def encoder_devices(self):
        return self.devices[1:-1]

    @property
    def head_device(self):
        return self.devices[-1]

    def log(self, message):
        if self.verbose:
            print(f"{datetime.now()} - {message}")

    def _module(self, idx):
        L = self.config.num_hidden_layers
        if idx == 0:
            return self.embeddings
        if idx == L + 1:
            return self.head
        return self.encoders[idx - 1]

    def to_pipeline(self, chunks: int, checkpoint: str = "never"):
        """GPipe over the same placement with `chunks` micro-batches (model/bert_mp.py:73-89).
        `checkpoint`: 'never' (default), 'except_last' (torch Pipe's default: recompute the forward of
        all micro-batches but the last during backward) or 'always'.  Recomputation only buys
        memory; at the reference's sizes a stage's activations are a few GB of 180, and on the last
        stage the extra forwards sit on the critical path of the whole pipeline."""
        self._chunks = max(1, int(chunks))
        self._checkpoint = checkpoint
        self._pipe = None
        return _Pipelined(self)

    # ------------------------------------------------------------------ single-process mode
    def _forward_local(self, input_ids):
        h = None
        for gi, rng in enumerate(self.group_ranges):
            tr = self.trackers[gi]
            dev = self.devices[rng[0]]
            tr.enter("forward pass")
            for idx in rng:
                m = self._module(idx)
                h = m(input_ids.to(dev)) if idx == 0 else m(h.to(dev))
            tr.leave()
        return h

    def forward(self, input_ids: torch.LongTensor):
        if self.distributed:
            raise RuntimeError("with one process per stage use forward_backward(input_ids, labels)")
        out = self._forward_local(input_ids)
        if out.requires_grad:
            out.register_hook(self._mark_backward_start)
        return out

    def _mark_backward_start(self, grad):
        # single-process backward runs through all groups in one autograd call; the per-device
        # segments cannot be separated from the host, so we time the whole backward per device
        return grad

    # ------------------------------------------------------------------ one process per stage
    def _stage(self) -> PipelineStage:
        if self._pipe is None or self._pipe.chunks != self._chunks:
            mods = [self._module(i) for i in self._mine]
            if self.stage == self.device_count - 1:
                mods = mods[:-1]  # the head runs inside the loss (fused projection + cross-entropy)
            head_params = list(self.head.parameters()) if self.stage == self.device_count - 1 else []
            self._pipe = PipelineStage(mods, self.stage, self.device_count, self.local_device,
                                       act_dtype=self.dtype, group=self.group, chunks=self._chunks,
                                       tracker=self.trackers[self.stage], extra_params=head_params,
                                       checkpoint=getattr(self, "_checkpoint", "never"))
        return self._pipe

    def forward_backward(self, input_ids, labels):
        """Runs forward and backward of one batch (GPipe if ``to_pipeline`` was called, else one
        micro-batch).  Returns the loss on the last stage, None elsewhere."""
        assert self.distributed
        st = self._stage()
        B, S = input_ids.shape

        def loss_fn(h, lab, n_valid):
            # sum of the micro-batch's token losses / labelled tokens of the whole batch
            return self.head(h, labels=lab, n_valid=n_valid)

        return st.forward_backward(input_ids, labels, (B, S, self.config.hidden_size), loss_fn)

    def collect_idle_time(self):
        """Fills ``device_idle_time[device] = (total_ms, count)`` for every device (gathered over
        ranks in distributed mode) and returns it."""
        local = {d: tr.totals() for d, tr in self.trackers.items()}
        if self.distributed:
            allv = [None] * self.device_count
            dist.all_gather_object(allv, local, group=self.group)
            for d in allv:
                for k, v in d.items():
                    self.device_idle_time[k] = v
        else:
            self.device_idle_time.update(local)
        return self.device_idle_time


class _Pipelined(nn.Module):
    """What ``to_pipeline`` returns.  In single-process mode calling it runs the micro-batches back
    to back and returns an object with ``.local_value()``; in distributed mode use
    ``forward_backward``."""

    def __init__(self, bert: BertModelWithMP):
        super().__init__()
        self.bert = bert

    def forward(self, input_ids):
        outs = [self.bert._forward_local(c) for c in input_ids.chunk(self.bert._chunks)]
        return LocalValue(torch.cat(outs, dim=0))

    def forward_backward(self, input_ids, labels):
        return self.bert.forward_backward(input_ids, labels)

    def parameters(self, recurse=True):
        return self.bert.parameters(recurse)

    def train(self, mode=True):
        self.bert.train(mode)
        return self
