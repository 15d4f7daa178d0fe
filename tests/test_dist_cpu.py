"""CPU / gloo world_size=2 plumbing tests (BASELINE.json config 1): process-group bootstrap,
logged collectives, DDP bucket schedule and gradient equivalence, training loop."""
import pytest

from _mp import run_distributed


def test_allreduce_demo_gloo_world3():
    run_distributed("allreduce_demo", world=3, backend="gloo")


def test_ddp_matches_single_process_gloo():
    run_distributed("ddp_matches_single_process", world=2, backend="gloo")


def test_ddp_training_loop_gloo():
    run_distributed("ddp_training_loop", world=2, backend="gloo", args=(False,))


def test_ddp_fused_optimizer_gloo():
    run_distributed("ddp_training_loop", world=2, backend="gloo", args=(False, True))


def test_optimizer_built_before_ddp_wrap_gloo():
    run_distributed("optimizer_built_before_ddp_wrap", world=2, backend="gloo")


def test_zero_stages_world3_gloo():
    """Bucket edges must split evenly over 3 ranks (8*world does not divide 64)."""
    run_distributed("zero_stages_match_stage0", world=3, backend="gloo", timeout=900)


def test_zero_stages_match_stage0_gloo():
    run_distributed("zero_stages_match_stage0", world=2, backend="gloo", timeout=600)


def test_gpipe_matches_single_process_gloo():
    run_distributed("pipeline_matches_single_process", world=2, backend="gloo", args=(4,), timeout=600)


def test_gpipe_recompute_replays_dropout_gloo():
    run_distributed("pipeline_recompute_replays_dropout", world=2, backend="gloo", args=(4,), timeout=600)


def test_zero_checkpoint_resume_gloo():
    run_distributed("zero_checkpoint_resume", world=2, backend="gloo", args=(2,), timeout=600)
    run_distributed("zero_checkpoint_resume", world=2, backend="gloo", args=(3,), timeout=600)


def test_pipeline_checkpoint_resume_gloo():
    run_distributed("pipeline_checkpoint_resume", world=2, backend="gloo", timeout=600)
