"""Multi-GPU tests (need >= 2 visible GPUs): custom P2P collectives vs NCCL, the fused ZeRO step
kernel, and DDP equivalence / training with and without CUDA graphs."""
import os

import pytest
import torch

from _mp import run_distributed

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")]


def test_allreduce_demo():
    run_distributed("allreduce_demo", world=2, backend="nccl")


def test_collectives_match_torch():
    run_distributed("collectives_match_torch", world=2, backend="nccl")


def test_zero_step_kernel():
    run_distributed("zero_step_matches_reference", world=2, backend="nccl")


def test_ddp_matches_single_process():
    run_distributed("ddp_matches_single_process", world=2, backend="nccl")


@pytest.mark.parametrize("use_graph", [False, True])
def test_ddp_training_loop(use_graph):
    run_distributed("ddp_training_loop", world=2, backend="nccl", args=(use_graph,))


@pytest.mark.parametrize("use_graph", [False, True])
def test_ddp_fused_optimizer(use_graph):
    run_distributed("ddp_training_loop", world=2, backend="nccl", args=(use_graph, True))


@pytest.mark.parametrize("arch", ["bloom-tiny", "bert-tiny"])
def test_zero_stages_match_stage0(arch):
    run_distributed("zero_stages_match_stage0", world=2, backend="nccl", args=(arch,), timeout=600)


@pytest.mark.parametrize("chunks", [1, 4])
def test_pipeline_matches_single_process(chunks):
    run_distributed("pipeline_matches_single_process", world=2, backend="nccl", args=(chunks,), timeout=600)


@pytest.mark.parametrize("checkpoint", ["never", "except_last"])
def test_pipeline_graphs_match_eager(checkpoint):
    run_distributed("pipeline_graphs_match_eager", world=2, backend="nccl", args=(4, checkpoint), timeout=600)


def test_pipeline_recompute_replays_dropout():
    run_distributed("pipeline_recompute_replays_dropout", world=2, backend="nccl", args=(4,), timeout=600)


def test_nvls_collectives_own_multicast_runtime():
    run_distributed("nvls_collectives", world=2, backend="nccl", args=("own-only",))


def test_optimizer_built_before_ddp_wrap():
    run_distributed("optimizer_built_before_ddp_wrap", world=2, backend="nccl")


def test_allgather_fused_gemm():
    run_distributed("allgather_gemm_matches_gathered", world=2, backend="nccl")


def test_ddp_nvls_gradient_allreduce(monkeypatch):
    """Same DDP equivalence with the gradient buffer bound to an NVSwitch multicast object and the
    buckets reduced by the multimem kernel (forced on: by default NVLS is used from 4 ranks up)."""
    monkeypatch.setenv("DTDS_NVLS", "1")
    run_distributed("ddp_matches_single_process", world=2, backend="nccl", args=(0.5,))
