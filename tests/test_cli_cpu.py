"""The reference's entry points, run end to end on CPU (gloo, world_size 2) with tiny shapes:
BASELINE.json config 1 (`model/transformer.py` block through `data_parallel_training.py`), the
ZeRO script under the in-repo launcher, the model-parallel script in pipeline mode."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, timeout=600, env=None):
    e = dict(os.environ)
    e.update(env or {})
    e["CUDA_VISIBLE_DEVICES"] = e.get("CUDA_VISIBLE_DEVICES", "")
    e.setdefault("OMP_NUM_THREADS", str(max(1, (os.cpu_count() or 2) // 2)))   # two ranks share the cores
    return subprocess.run([sys.executable, *args], cwd=ROOT, capture_output=True, text=True, timeout=timeout, env=e)


def test_ddp_script_transformer_block_gloo_world2():
    r = _run(["data_parallel_training.py", "--backend", "gloo", "--device-count", "2", "--model", "block",
              "--batch-size", "2", "--training-steps", "8", "--seq-len", "64", "--port", "12421"])
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.count("Total Training Time:") == 2   # every rank prints it, like the reference


def test_zero_script_under_launcher_gloo_world2():
    for stage in (2, 3):
        r = _run(["-m", "dtds_b200.launch", "--nproc-per-node", "2", "zero_dp_training.py", "--stage", str(stage),
                  "--model_name", "bloom-tiny", "--training-steps", "6", "--seq-len", "64", "--quiet"])
        out = r.stdout + r.stderr
        assert r.returncode == 0, out[-3000:]
        assert f"ZeRO Stage: {stage}" in out and "Optimizer: lr=0.00015; betas=(0.9, 0.999); eps=1e-08" in out
        assert "Total Training Time:" in out and "Total Communication Latency:" in out
        assert "Comm. Op" in out   # the summary table


def test_model_parallel_script_pipeline_gloo_world2():
    r = _run(["model_parallel_training.py", "--backend", "gloo", "--device-count", "2", "--model", "tiny",
              "--batch-size", "8", "--training-steps", "3", "--seq-len", "64", "--pipeline", "--port", "12423"])
    out = r.stdout + r.stderr
    assert r.returncode == 0, out[-3000:]
    assert "Total Training Time:" in out and "Average Idle Time (ms)" in out


def test_scripts_save_and_resume_with_real_text(tmp_path):
    """--data (local text), --save-dir and --resume through the DDP and ZeRO entry points."""
    text = tmp_path / "corpus.txt"
    text.write_text("\n".join(f"Passage {i} about topic {i % 7} : " + " ".join(f"w{(i + j) % 40}" for j in range(30))
                              for i in range(64)))
    ck = tmp_path / "ddp_ckpt"
    common = ["data_parallel_training.py", "--backend", "gloo", "--device-count", "2", "--model", "tiny",
              "--batch-size", "2", "--seq-len", "64", "--data", str(text)]
    r = _run(common + ["--training-steps", "4", "--port", "12431", "--save-dir", str(ck)])
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert sorted(os.listdir(ck)) == ["meta.pt", "model.pt", "optim_rank0.pt", "optim_rank1.pt"]
    r = _run(common + ["--training-steps", "6", "--port", "12432", "--resume", str(ck)])
    assert r.returncode == 0 and "Resumed from" in r.stdout and "at step 2" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
    zk = tmp_path / "zero_ckpt"
    zero = ["-m", "dtds_b200.launch", "--nproc-per-node", "2", "zero_dp_training.py", "--stage", "3", "--model-name",
            "opt-tiny", "--seq-len", "64", "--quiet", "--data", str(text)]
    r = _run(zero + ["--training-steps", "4", "--save-dir", str(zk)])
    assert r.returncode == 0 and "Saved checkpoint" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
    r = _run(zero + ["--training-steps", "6", "--resume", str(zk)])
    assert r.returncode == 0 and "Resumed from" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
