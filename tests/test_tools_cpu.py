"""Golden values of the estimator tools, util helpers, synthetic data, comms logger, launcher,
bucket assignment and checkpoint resume — everything that needs no GPU and no process group."""
import os
import subprocess
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args, timeout=300):
    return subprocess.run([sys.executable, *args], cwd=ROOT, capture_output=True, text=True, timeout=timeout)


def test_estimate_nn_memory_golden_values():
    r = _run("estimate_nn_memory.py")
    assert "Model Memory: 12,597,248 bytes" in r.stdout, r.stdout + r.stderr
    assert "Activation Memory: 65,536 bytes" in r.stdout


def test_transformer_memory_projection_golden_value():
    sys.path.insert(0, ROOT)
    import estimate_transformer_memory as etm
    total = etm.project_transformer_memory(1, 9216, 72, 4, 512, "adam")
    assert int(total) == 18_253_430_784
    b = etm.project_transformer_memory(1, 9216, 72, 4, 512, "adam", breakdown=True)
    assert (int(b["model"]), int(b["activations"]), int(b["optimizer"])) == (4_077_342_720, 1_944_059_904, 8_154_685_440)
    from util import format_size
    assert format_size(total) == "17.0 GB"
    fused = etm.project_transformer_memory(1, 9216, 72, 4, 512, "adam", fused_attention=True)
    assert fused < total
    r = _run("estimate_transformer_memory.py", "--no-measure")
    assert "Projected total memory usage: 17.0 GB" in r.stdout, r.stdout + r.stderr


def test_util_and_synthetic_dataset():
    sys.path.insert(0, ROOT)
    from util import format_size, get_device, get_device_count, load_wikitext, synthetic_collator, synthetic_tokenizer
    from dtds_b200.models import get_config
    assert format_size(0) == "0B" and format_size(1536) == "1.5 KB"
    assert get_device() in ("cuda", "cpu", "mps") and get_device_count() >= 1
    tok = synthetic_tokenizer(get_config("bert-base-cased"))
    ds = load_wikitext(tok, synthetic_collator(tok, mlm=True)).select(range(32))
    assert len(ds) == 32 and ds[0]["input_ids"].shape == (512,) and ds[0]["labels"].dtype == torch.int64
    frac = (ds.labels != -100).float().mean().item()
    assert 0.10 < frac < 0.20
    assert (ds.input_ids[ds.labels != -100] == 103).float().mean().item() > 0.6   # ~80 % [MASK]
    clm = load_wikitext(tok, synthetic_collator(tok, mlm=False), max_length=128).select(range(4))
    assert torch.equal(clm.input_ids, clm.labels) and clm.input_ids.shape == (4, 128)


def test_comms_logger_data_model_and_busbw():
    from dtds_b200.comm.logger import CommsLogger, calc_bw
    lg = CommsLogger()
    lg.configure({"enabled": True, "verbose": False, "prof_all": True, "debug": False})
    lg.append_host("all_reduce", 1 << 20, 1.0, 8)
    lg.append_host("all_reduce", 1 << 20, 3.0, 8)
    rec = lg.comms_dict["all_reduce"][1 << 20]
    assert rec[0] == 2 and rec[1] == [1.0, 3.0] and len(rec[2]) == 2 and len(rec[3]) == 2
    alg, bus = calc_bw("all_reduce", 10 ** 9, 1000.0, 8)
    assert abs(alg - 8.0) < 1e-9 and abs(bus - 8.0 * 2 * 7 / 8) < 1e-9
    alg, bus = calc_bw("all_gather_into_tensor", 10 ** 9, 1000.0, 4)
    assert abs(bus - 8.0 * 3 / 4) < 1e-9
    assert lg.total_latency_ms() == 4.0
    lg.log_all(print_log=True)


def test_bucket_assignment_reverse_order_and_caps():
    from dtds_b200.parallel.ddp import assign_buckets
    ps = [torch.nn.Parameter(torch.zeros(n)) for n in (300_000, 100_000, 100_000, 200_000, 50_000)]
    offs, o = [], 0
    for p in ps:
        offs.append(o)
        o += p.numel()
    b = assign_buckets(ps, offs, bucket_cap_mb=1.0, elem_size=4, first_bucket_mb=0.25)  # caps: 262144 / 65536 elts
    assert b[0][0] == [4]                     # small first bucket holds only the last parameter
    # once no more than one cap's worth of ordinary parameters is left (the two 100k ones), the cap
    # drops to half of that remainder: they are reduced separately; the oversize parameter sits alone
    assert b[1][0] == [3] and b[2][0] == [2] and b[3][0] == [1] and b[4][0] == [0]
    assert all(lo < hi for _, lo, hi in b)


def test_launcher_env_contract_and_failure_propagation(tmp_path):
    script = tmp_path / "child.py"
    script.write_text("import os, sys\n"
                      "assert any(a.startswith('--local_rank=') for a in sys.argv)\n"
                      # one write() per line: the two unbuffered children share the pipe
                      "os.write(1, ('R %s %s %s %s\\n' % (os.environ['RANK'], os.environ['LOCAL_RANK'], "
                      "os.environ['WORLD_SIZE'], os.environ['MASTER_ADDR'])).encode())\n"
                      "sys.exit(3 if os.environ['RANK'] == '1' and 'fail' in sys.argv else 0)\n")
    ok = _run("-m", "dtds_b200.launch", "--nproc-per-node", "2", str(script))
    assert ok.returncode == 0 and "R 0 0 2 127.0.0.1" in ok.stdout and "R 1 1 2 127.0.0.1" in ok.stdout
    bad = _run("-m", "dtds_b200.launch", "--nproc-per-node", "2", str(script), "fail")
    assert bad.returncode == 3
    from dtds_b200.launch import parse_hostfile
    hf = tmp_path / "hostfile"
    hf.write_text("worker-1 slots=4\nworker-2 slots=2  # comment\n")
    assert parse_hostfile(str(hf)) == {"worker-1": 4, "worker-2": 2}


def test_checkpoint_resume_matches_uninterrupted_run(tmp_path):
    from dtds_b200.models import build_model, get_config
    from dtds_b200.optim import FusedAdam
    from dtds_b200.utils.ckpt import load_checkpoint, save_checkpoint
    cfg = get_config("opt-tiny").replace(hidden_dropout_prob=0.0, num_hidden_layers=1)
    ids = torch.randint(0, cfg.vocab_size, (2, 16), generator=torch.Generator().manual_seed(0))

    def make():
        torch.manual_seed(0)
        m = build_model(cfg)
        return m, FusedAdam(m.parameters(), lr=1e-2)

    def steps(m, opt, n):
        out = []
        for _ in range(n):
            opt.zero_grad()
            loss = m(ids, labels=ids).loss
            loss.backward()
            opt.step()
            out.append(loss.item())
        return out

    m1, o1 = make()
    full = steps(m1, o1, 6)
    m2, o2 = make()
    first = steps(m2, o2, 3)
    save_checkpoint(str(tmp_path / "ck"), m2, o2, step=3)
    m3, o3 = make()
    with torch.no_grad():
        for p in m3.parameters():
            p.add_(1.0)   # scramble: the load must restore everything
    assert load_checkpoint(str(tmp_path / "ck"), m3, o3) == 3
    rest = steps(m3, o3, 3)
    assert first + rest == full
    sd = torch.load(tmp_path / "ck" / "model.pt")
    assert set(sd.keys()) == set(m1.state_dict().keys())


def test_nvtx_ranges_are_balanced_and_free_when_off(monkeypatch):
    """DTDS_NVTX ranges: nothing is emitted when off; when on, every push has its pop across the
    model annotation, the DDP bucket launches and the optimizer step."""
    from dtds_b200.models import build_model, get_config
    from dtds_b200.optim import FusedAdam
    from dtds_b200.parallel.ddp import DistributedDataParallel
    from dtds_b200.utils import trace
    events = []
    monkeypatch.setattr(trace, "_push", lambda name: events.append(("push", name)))
    monkeypatch.setattr(trace, "_pop", lambda: events.append(("pop", None)))
    torch.manual_seed(0)
    m = build_model(get_config("bert-tiny"))
    ddp = DistributedDataParallel(m, bucket_cap_mb=0.1)
    opt = FusedAdam(m.parameters(), lr=1e-3)
    handles = trace.annotate_model(m)
    ids = torch.randint(0, 100, (2, 128))

    def step():
        with trace.nvtx_range("train step"):
            opt.zero_grad()
            ddp(ids, labels=ids).loss.backward()
            opt.step()

    trace.set_nvtx(False)
    step()
    assert events == []
    trace.set_nvtx(True)
    try:
        step()
    finally:
        trace.set_nvtx(False)
        for h in handles:
            h.remove()
    depth = 0
    for kind, _ in events:
        depth += 1 if kind == "push" else -1
        assert depth >= 0
    assert depth == 0
    names = [n for k, n in events if k == "push"]
    assert names[0] == "train step" and "FusedAdam.step" in names
    assert any(n.startswith("encoders.0 ") for n in names) and any(n.startswith("ddp.bucket") for n in names)
    t = trace.DeviceTimer("cpu")
    with t:
        sum(range(1000))
    assert t.count == 1 and t.ms(reduce=False) > 0


def test_real_text_dataset_follows_the_reference_recipe(tmp_path):
    """Local text file -> rows shaped like util.load_wikitext's: padded to max_length, static
    15 % MLM masks on real tokens only, -100 elsewhere; causal rows copy input_ids."""
    from dtds_b200.models import get_config
    from dtds_b200.utils.data import load_wikitext, synthetic_collator, synthetic_tokenizer
    text = tmp_path / "wiki.train.raw"
    lines = ["", " = Valkyria Chronicles III = ", ""] + [
        f"Passage {i} : " + " ".join(f"word{(i * 7 + j) % 50}" for j in range(20 + i % 30)) + " ." for i in range(200)]
    text.write_text("\n".join(lines))
    cfg = get_config("bert-tiny")
    tok = synthetic_tokenizer(cfg)
    mlm = load_wikitext(tok, synthetic_collator(tok, mlm=True), max_length=64, path=str(text))
    assert len(mlm) == 201 and mlm[0]["input_ids"].shape == (64,)
    ids, labels, real = mlm.input_ids, mlm.labels, mlm.attention_mask
    assert (ids[~real] == 0).all() and (labels[~real] == -100).all()
    frac = (labels[real] != -100).float().mean().item()
    assert 0.11 < frac < 0.19, frac
    picked = labels != -100
    assert 0.7 < (ids[picked] == tok.mask_token_id).float().mean().item() < 0.9
    assert ids.max().item() < cfg.vocab_size and ids[real & ~picked].min().item() >= 500
    again = load_wikitext(tok, synthetic_collator(tok, mlm=True), max_length=64, path=str(text))
    assert torch.equal(again.input_ids, ids)                      # static masks, deterministic
    clm = load_wikitext(tok, synthetic_collator(tok, mlm=False), max_length=64, path=str(text))
    assert torch.equal(clm.labels[clm.attention_mask], clm.input_ids[clm.attention_mask])
    assert (clm.labels[~clm.attention_mask] == -100).all()
    sub = mlm.select(range(300))                                   # more rows than lines: wraps
    assert len(sub) == 300 and torch.equal(sub.input_ids[201], mlm.input_ids[0])


def test_launcher_multi_node_env_and_ssh_fanout(tmp_path):
    """Two "nodes" on this machine: ranks are node_rank * slots + local_rank, every node sees the
    same MASTER_ADDR/PORT, a gloo all-reduce crosses the nodes, and comm reports that the job spans
    boxes (which keeps the engines off the peer-memory kernels).  The ssh fan-out is driven through
    a stand-in for ssh that runs the remote command locally."""
    child = tmp_path / "child.py"
    child.write_text(
        "import os, sys\n"
        f"sys.path.insert(0, {ROOT!r})\n"
        "import torch\n"
        "from dtds_b200 import comm\n"
        "comm.init_process_group(backend='gloo')\n"
        "t = torch.tensor([float(comm.get_rank() + 1)])\n"
        "comm.all_reduce(t)\n"
        "os.write(1, ('R%s/%s local=%s/%s node=%s sum=%.0f spans=%s\\n' % (os.environ['RANK'], os.environ['WORLD_SIZE'],"
        " os.environ['LOCAL_RANK'], os.environ['LOCAL_WORLD_SIZE'], os.environ['NODE_RANK'], t.item(),"
        " comm.spans_nodes())).encode())\n"
        "comm.destroy_process_group()\n")
    fake_ssh = tmp_path / "fake_ssh.sh"
    fake_ssh.write_text('#!/bin/sh\n# usage: fake_ssh.sh HOST COMMAND\nshift\nexec sh -c "$1"\n')
    fake_ssh.chmod(0o755)
    hostfile = tmp_path / "hostfile"
    hostfile.write_text("127.0.0.1 slots=2\nlocalhost slots=2\n")
    from dtds_b200.launch import free_port
    r = _run("-m", "dtds_b200.launch", "--hostfile", str(hostfile), "--ssh", str(fake_ssh),
             "--master-port", str(free_port()), "--no-local-rank", str(child))
    assert r.returncode == 0, r.stdout + r.stderr
    lines = sorted(l for l in r.stdout.splitlines() if l.startswith("R"))
    assert lines == ["R0/4 local=0/2 node=0 sum=10 spans=True", "R1/4 local=1/2 node=0 sum=10 spans=True",
                     "R2/4 local=0/2 node=1 sum=10 spans=True", "R3/4 local=1/2 node=1 sum=10 spans=True"], r.stdout


def test_cluster_templates_render_hostfile_and_ssh_config():
    """aws/cluster.py is the pure-Python half of the CDK stacks (the stacks themselves need
    aws-cdk-lib): its hostfile is what the launcher parses."""
    sys.path.insert(0, os.path.join(ROOT, "aws"))
    try:
        import cluster
        import training_stack  # noqa: F401  (importable without aws-cdk-lib)
    finally:
        sys.path.pop(0)
    from dtds_b200.launch import parse_hostfile
    names = ["ec2-1.compute.amazonaws.com", "ec2-2.compute.amazonaws.com"]
    hf = cluster.render_hostfile(names)
    import tempfile
    with tempfile.NamedTemporaryFile("w", suffix="hostfile", delete=False) as f:
        f.write(hf)
    assert parse_hostfile(f.name) == {names[0]: 8, names[1]: 8}
    cfg = cluster.render_ssh_config(names, key_file="~/k.pem")
    assert cfg.count("Host worker-") == 2 and "HostName ec2-2.compute.amazonaws.com" in cfg and "IdentityFile ~/k.pem" in cfg
    spec = cluster.ClusterSpec(nodes=3)
    assert spec.worker_names() == ["worker-1", "worker-2", "worker-3"] and spec.instance_type.startswith("p6-b200")


def test_native_text_packer_matches_python_tokenizer(tmp_path):
    """csrc/text_rows.cpp (multi-threaded file -> padded token rows) against the Python statement
    of the same byte-level rules, on text with punctuation, tabs, upper case and UTF-8."""
    from dtds_b200.utils.data import WordHashTokenizer
    tok = WordHashTokenizer(vocab_size=28996, model_max_length=48)
    lines = ["Hello, World!  It's 2024 …", "", "\t  ", "naïve café — Ünïcödé_words and snake_case_9",
             "x" * 300, "a b c " * 40] + [f"line {i}: The QUICK brown-fox (#{i}) jumps; over {i * 17} dogs." for i in range(600)]
    path = tmp_path / "text.txt"
    path.write_text("\n".join(lines), encoding="utf-8")
    packed = tok.encode_file(str(path), 48)
    assert packed is not None, "native library not built"
    ids, lens = packed
    want = [tok.encode(l)[:48] for l in lines if l.strip()]
    assert ids.shape == (len(want), 48)
    for r, w in enumerate(want):
        assert lens[r].item() == len(w)
        assert ids[r, :len(w)].tolist() == w and (ids[r, len(w):] == tok.pad_token_id).all()
    assert ids.min().item() >= 0 and ids[ids != 0].min().item() >= tok.first_id and ids.max().item() < 28996
