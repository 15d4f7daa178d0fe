"""Whole-step CUDA-graph capture.

At the reference's batch sizes (4 x 512 tokens per rank for DDP, 1 x 512 for ZeRO) a training
step is hundreds of short kernels; launched eagerly from Python the GPU idles between them.
``GraphedStep`` captures forward + backward (including the side-stream bucket collectives the
engines launch from gradient-ready hooks) + the optimizer into ONE CUDA graph and replays it per
step: no Python, no launch gaps, static addresses.  This is the "CUDA streams and graphs instead
of a tracing compiler" execution model of the framework.

What makes the step capturable:
  * no host synchronisation anywhere in the step (losses stay on the device; the Adam step count
    is a device scalar; cross-entropy normalisation is computed on the device);
  * dropout seeds are fixed at capture time but every kernel mixes in a device-side offset that
    the graph itself increments, so each replay draws fresh masks;
  * inputs are copied into static buffers (one H2D copy from pinned memory per step).
"""
from __future__ import annotations

import torch

from ..ops import kernels as K


class GraphedStep:
    """step_fn(*static_inputs) -> loss tensor, captured once.

    ``step_fn`` must run one full training step (zero_grad, forward, backward, optimizer step)
    and return the scalar loss.  ``example_inputs`` fix shapes and dtypes."""

    def __init__(self, step_fn, example_inputs, warmup: int = 3, device=None):
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.step_fn = step_fn
        self.static_inputs = [torch.empty_like(t, device=self.device) for t in example_inputs]
        for s, t in zip(self.static_inputs, example_inputs):
            s.copy_(t)
        self.seed_offset = torch.zeros(1, dtype=torch.int64, device=self.device)
        self.graph = None
        self.static_loss = None
        self._stream = torch.cuda.Stream(self.device)
        # warm-up on a side stream (lazy allocations, cudaFuncSetAttribute calls, autotuning)
        self._stream.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(self._stream):
            for _ in range(warmup):
                self._one()
        torch.cuda.current_stream(self.device).wait_stream(self._stream)
        torch.cuda.synchronize(self.device)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, stream=self._stream):
            self.static_loss = self._one()
        torch.cuda.synchronize(self.device)

    def _one(self):
        # The device-side seed offset is only registered while this step is being recorded (its
        # address gets baked into the captured kernels); eager code elsewhere is unaffected.
        K.set_seed_offset_tensor(self.seed_offset)
        try:
            loss = self.step_fn(*self.static_inputs)
        finally:
            K.set_seed_offset_tensor(None)
        self.seed_offset.add_(1)
        return loss

    def load_inputs(self, *inputs, non_blocking: bool = True):
        for s, t in zip(self.static_inputs, inputs):
            s.copy_(t, non_blocking=non_blocking)

    def replay(self) -> torch.Tensor:
        self.graph.replay()
        return self.static_loss

    def __call__(self, *inputs) -> torch.Tensor:
        self.load_inputs(*inputs)
        return self.replay()
