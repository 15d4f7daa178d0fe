"""CPU checks of the model families: parameter counts of SURVEY.md §2.6, state-dict names,
equivalence with HF BERT, TransformerBlock parity with an HF OPT layer, config surface."""
import pytest
import torch

from dtds_b200.models import (BertForMaskedLM, TransformerBlock, build_model, count_parameters,
                              from_hf_state_dict, get_config, to_hf_state_dict)
from dtds_b200.models.bert_mp import BertModelWithMP
from dtds_b200.parallel.pipeline import split_sizes


@pytest.mark.parametrize("name,count", [("bert-base-cased", 108_340_804), ("bert-large-cased", 333_610_308),
                                        ("bigscience/bloom-560m", 559_214_592), ("facebook/opt-125m", 125_239_296)])
def test_parameter_counts(name, count):
    with torch.device("meta"):
        model = build_model(get_config(name))
    assert count_parameters(model) == count


def test_bert_mp_untied_parameter_count_and_prefixes():
    cfg = get_config("bert-base-cased")
    with torch.device("meta"):
        from dtds_b200.models.bert import BertEmbeddings, BertLayer, BertOnlyMLMHead
        u = cfg.replace(tie_word_embeddings=False)
        n = sum(p.numel() for p in BertEmbeddings(u).parameters())
        n += 12 * sum(p.numel() for p in BertLayer(u).parameters())
        n += sum(p.numel() for p in BertOnlyMLMHead(u).parameters())
    assert n == 130_609_732
    tiny = BertModelWithMP(get_config("bert-tiny"), device_count=1)
    keys = list(tiny.state_dict().keys())
    assert all(k.split(".")[0] in ("embeddings", "encoders", "head") for k in keys)
    assert any(k.startswith("encoders.1.") for k in keys)


def test_array_split_group_sizes():
    assert split_sizes(14, 2) == [7, 7]
    assert split_sizes(14, 3) == [5, 5, 4]
    assert split_sizes(14, 4) == [4, 4, 3, 3]
    assert split_sizes(14, 8) == [2, 2, 2, 2, 2, 2, 1, 1]


def test_bert_matches_hf_implementation():
    """Same weights (via the HF-name bridge) -> same logits and same loss as transformers' BERT."""
    transformers = pytest.importorskip("transformers")
    cfg = get_config("bert-tiny").replace(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    hf_cfg = transformers.BertConfig(vocab_size=cfg.vocab_size, hidden_size=cfg.hidden_size,
                                     num_hidden_layers=cfg.num_hidden_layers,
                                     num_attention_heads=cfg.num_attention_heads,
                                     intermediate_size=cfg.intermediate_size,
                                     max_position_embeddings=cfg.max_position_embeddings,
                                     hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    torch.manual_seed(0)
    hf = transformers.BertForMaskedLM(hf_cfg).eval()
    mine = BertForMaskedLM(cfg).eval()
    from_hf_state_dict(mine, hf.state_dict())
    ids = torch.randint(0, cfg.vocab_size, (2, 64))
    labels = ids.clone()
    labels[:, ::2] = -100
    with torch.no_grad():
        ref = hf(input_ids=ids, labels=labels)
        got = mine(ids, labels=labels, return_logits=True)
    assert torch.allclose(got.logits, ref.logits, atol=2e-4, rtol=1e-4)
    assert abs(got.loss.item() - ref.loss.item()) < 1e-4
    # and back: exported names load into the HF model
    exported = to_hf_state_dict(mine)
    missing, unexpected = hf.load_state_dict(exported, strict=False)
    assert not [k for k in unexpected], unexpected


def test_transformer_block_names_and_param_count():
    from types import SimpleNamespace
    cfg = SimpleNamespace(hidden_size=128, num_attention_heads=2, ffn_dim=512, enable_bias=True, dropout=0.1,
                          max_position_embeddings=64)
    blk = TransformerBlock(cfg)
    names = {n.split(".")[0] for n, _ in blk.named_parameters()}
    assert names == {"pre_attention_layer_norm", "k_proj", "v_proj", "q_proj", "out_proj",
                     "final_layer_norm", "ffn_1", "ffn_2"}
    mods = {n for n, _ in blk.named_children()}
    assert {"compute_attention_weights", "attention_weights_softmax", "attention_weights_dropout",
            "compute_attentions", "attention_output_dropout", "activation", "final_dropout"} <= mods
    h = 128
    assert sum(p.numel() for p in blk.parameters()) == 12 * h * h + 13 * h
    # h = 9216 closed form of the estimator (1,019,335,680)
    assert 12 * 9216 * 9216 + 13 * 9216 == 1_019_335_680
    x = torch.randn(2, 16, h)
    blk.eval()
    assert blk(x).shape == x.shape


def test_causal_lm_shift_and_tied_head():
    cfg = get_config("opt-tiny").replace(hidden_dropout_prob=0.0)
    m = build_model(cfg).eval()
    ids = torch.randint(0, cfg.vocab_size, (2, 32))
    out = m(ids, labels=ids, return_logits=True)
    ref = torch.nn.functional.cross_entropy(out.logits[:, :-1].reshape(-1, cfg.vocab_size), ids[:, 1:].reshape(-1))
    assert abs(out.loss.item() - ref.item()) < 1e-4
    assert m.head_weight is m.word_embeddings


def _train_three_steps(overlap=False, ddp=False, opt_first=False, bucket_mb=0.05, deep=False):
    """Three AdamW steps on bert-tiny (eval mode: no dropout noise); returns the flat weights."""
    from dtds_b200.optim import HFAdamW
    from dtds_b200.parallel.ddp import DistributedDataParallel as DDP
    torch.manual_seed(0)
    m = build_model(get_config("bert-tiny"))
    if deep:   # deepcopy drops Python attributes of Parameters (the tied-embedding mark)
        import copy
        m = copy.deepcopy(m)
    m.eval()
    opt = None
    if opt_first:   # the reference builds the optimizer before wrapping (data_parallel_training.py:34-44)
        opt = HFAdamW(m.parameters(), lr=1e-3)
    net = DDP(m, bucket_cap_mb=bucket_mb) if ddp else m
    if opt is None:
        opt = HFAdamW(m.parameters(), lr=1e-3, overlap_with_backward=overlap and not ddp,
                      overlap_bucket_mb=bucket_mb)
    if ddp and overlap:
        net.fuse_optimizer(opt)
    g = torch.Generator().manual_seed(1)
    for _ in range(3):
        ids = torch.randint(0, m.cfg.vocab_size, (2, 128), generator=g)
        opt.zero_grad()
        net(input_ids=ids, labels=ids).loss.backward()
        opt.step()
    return torch.cat([p.detach().flatten() for p in m.parameters()]), opt


def test_adam_overlapped_with_backward_is_exact():
    """Per-bucket Adam fired from the gradient-ready callbacks (own buckets, or the DDP engine's)
    gives bit-identical weights to one update after backward."""
    base, _ = _train_three_steps()
    own, opt = _train_three_steps(overlap=True)
    assert len(opt._obuckets) > 5
    assert torch.equal(base, own)
    fused, _ = _train_three_steps(overlap=True, ddp=True)
    assert torch.equal(base, fused)
    copied, _ = _train_three_steps(overlap=True, deep=True)
    assert torch.equal(base, copied)


def test_optimizer_built_before_ddp_wrap_follows_parameters():
    base, _ = _train_three_steps()
    late, opt = _train_three_steps(ddp=True, opt_first=True)
    assert torch.equal(base, late)
    assert getattr(opt.flats[0], "_superseded_by", None) is None


@pytest.mark.parametrize("name", ["bert-tiny", "bloom-tiny", "opt-tiny"])
def test_fused_skip_gradients_match_autograd_sum(name):
    """``fork=True`` outputs route the skip-branch gradient into the dgrad GEMM epilogue / the
    LayerNorm backward; with the switch off autograd sums the two branches itself."""
    from dtds_b200.ops import functional as Fn
    torch.manual_seed(0)
    cfg = get_config(name).replace(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    m = build_model(cfg)
    ids = torch.randint(0, cfg.vocab_size, (2, 128))
    grads = []
    for fused in (True, False):
        old = Fn.FUSE_SKIP_GRADS
        Fn.FUSE_SKIP_GRADS = fused
        try:
            for p in m.parameters():
                p.grad = None
            m(ids, labels=ids).loss.backward()
        finally:
            Fn.FUSE_SKIP_GRADS = old
        grads.append(torch.cat([p.grad.flatten() for p in m.parameters() if p.grad is not None]))
    assert grads[0].shape == grads[1].shape and grads[0].abs().sum() > 0
    torch.testing.assert_close(grads[0], grads[1], rtol=1e-4, atol=1e-6)


def _hf_causal(name):
    import transformers
    cfg = get_config(name)
    if cfg.arch == "bloom":
        hc = transformers.BloomConfig(vocab_size=cfg.vocab_size, hidden_size=cfg.hidden_size,
                                      n_layer=cfg.num_hidden_layers, n_head=cfg.num_attention_heads,
                                      layer_norm_epsilon=cfg.layer_norm_eps, hidden_dropout=0.0, attention_dropout=0.0)
        return cfg, transformers.BloomForCausalLM(hc).eval()
    hc = transformers.OPTConfig(vocab_size=cfg.vocab_size, hidden_size=cfg.hidden_size,
                                num_hidden_layers=cfg.num_hidden_layers, ffn_dim=cfg.intermediate_size,
                                num_attention_heads=cfg.num_attention_heads,
                                max_position_embeddings=cfg.max_position_embeddings,
                                word_embed_proj_dim=cfg.hidden_size, do_layer_norm_before=True,
                                activation_function="relu", dropout=0.0)
    return cfg, transformers.OPTForCausalLM(hc).eval()


@pytest.mark.parametrize("name", ["bloom-tiny", "opt-tiny"])
def test_causal_lm_matches_hf_implementation(name):
    """Same weights -> same logits, loss and gradients as HF BloomForCausalLM / OPTForCausalLM (the
    models zero_dp_training.py:24 loads), through the state-dict bridge in both directions."""
    from dtds_b200.models import causal_lm_from_hf_state_dict, causal_lm_to_hf_state_dict
    torch.manual_seed(0)
    cfg, hf = _hf_causal(name)
    mine = build_model(cfg).eval()
    causal_lm_from_hf_state_dict(mine, hf.state_dict())
    ids = torch.randint(0, cfg.vocab_size, (2, 128))
    a = hf(input_ids=ids, labels=ids)
    b = mine(ids, labels=ids, return_logits=True)
    torch.testing.assert_close(b.logits, a.logits, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(b.loss, a.loss, rtol=1e-5, atol=1e-5)
    a.loss.backward()
    mine(ids, labels=ids).loss.backward()
    g_hf = hf.get_input_embeddings().weight.grad
    torch.testing.assert_close(mine.word_embeddings.grad, g_hf, rtol=1e-3, atol=1e-6)
    # and back: a fresh HF model loaded from our export computes the same logits
    _, hf2 = _hf_causal(name)
    missing, unexpected = hf2.load_state_dict(causal_lm_to_hf_state_dict(mine), strict=False)
    assert not unexpected and all("masked_bias" in k or "attn.bias" in k for k in missing), (missing, unexpected)
    with torch.no_grad():
        torch.testing.assert_close(hf2(input_ids=ids).logits, a.logits.detach(), rtol=1e-4, atol=1e-4)


def test_load_pretrained_from_local_hf_checkpoint(tmp_path):
    """The offline stand-in for ``from_pretrained``: a HF state dict on disk goes through the bridge."""
    from dtds_b200.models import load_pretrained
    torch.manual_seed(0)
    cfg, hf = _hf_causal("opt-tiny")
    torch.save(hf.state_dict(), tmp_path / "pytorch_model.bin")
    mine = load_pretrained(build_model(cfg).eval(), str(tmp_path))
    ids = torch.randint(0, cfg.vocab_size, (1, 128))
    with torch.no_grad():
        torch.testing.assert_close(mine(ids, return_logits=True).logits, hf(input_ids=ids).logits,
                                   rtol=1e-4, atol=1e-4)


def test_clip_grad_norm_on_flat_gradients():
    """Gradients written straight into the flat buffer (p.grad stays None) are clipped like
    torch.nn.utils.clip_grad_norm_ clips ordinary ones."""
    from dtds_b200.optim import FusedAdam, clip_grad_norm_
    torch.manual_seed(0)
    cfg = get_config("bert-tiny").replace(hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    a = build_model(cfg)
    b = build_model(cfg)
    b.load_state_dict(a.state_dict())
    ids = torch.randint(0, cfg.vocab_size, (2, 128))
    opt = FusedAdam(a.parameters(), lr=1e-3)          # moves a's gradients into a flat buffer
    opt.zero_grad()
    a(ids, labels=ids).loss.backward()
    b(ids, labels=ids).loss.backward()
    assert all(p.grad is None for p in a.parameters())
    want = torch.nn.utils.clip_grad_norm_(b.parameters(), 0.5)
    got = clip_grad_norm_(opt, 0.5)
    torch.testing.assert_close(got, want, rtol=1e-4, atol=1e-6)
    for pa, pb in zip(a.parameters(), b.parameters()):
        torch.testing.assert_close(pa.main_grad, pb.grad, rtol=1e-4, atol=1e-7)


def test_zero_engine_rejects_ds_config_keys_it_does_not_implement():
    from dtds_b200.parallel import zero
    m = build_model(get_config("bloom-tiny"))
    base = {"train_micro_batch_size_per_gpu": 1, "optimizer": {"type": "Adam", "params": {"lr": 1e-3}},
            "zero_optimization": {"stage": 0}}
    for extra in ({"gradient_accumulation_steps": 4}, {"gradient_clipping": 1.0}, {"fp16": {"enabled": True}},
                  {"zero_optimization": {"stage": 2, "offload_optimizer": {"device": "cpu"}}}):
        with pytest.raises(NotImplementedError):
            zero.initialize(model=m, model_parameters=m.parameters(), config={**base, **extra})
    eng, opt, _, _ = zero.initialize(model=m, model_parameters=m.parameters(),
                                     config={**base, "gradient_accumulation_steps": 1, "gradient_clipping": 0.0})
    assert eng.zero_optimization_stage() == 0


def test_checkpoint_replays_fused_dropout_seed_stream():
    """ops.functional.checkpoint: recomputation regenerates the masks of the first forward (the seed
    stream is replayed) and leaves the stream where other consumers moved it in the meantime."""
    import torch
    from dtds_b200.models import get_config
    from dtds_b200.models.bert import BertLayer
    from dtds_b200.ops import functional as Fn
    cfg = get_config("bert-tiny").replace(hidden_dropout_prob=0.3, attention_probs_dropout_prob=0.3)
    torch.manual_seed(0)
    layers = torch.nn.ModuleList([BertLayer(cfg) for _ in range(2)]).train()
    x0 = torch.randn(2, 128, cfg.hidden_size)

    def run(h):
        for m in layers:
            h = m(h)
        return h

    res = {}
    for mode in ("plain", "checkpoint"):
        for p in layers.parameters():
            p.grad = None
        Fn.manual_seed(5)
        x = x0.clone().requires_grad_(True)
        y = run(x) if mode == "plain" else Fn.checkpoint(run, x)
        mark = Fn.next_seed()                 # another consumer between forward and backward
        (y ** 2).mean().backward()
        after = Fn.get_seed_state()
        res[mode] = (y.detach(), x.grad, [p.grad.clone() for p in layers.parameters()], mark, after)
    assert res["plain"][3] == res["checkpoint"][3] and res["plain"][4] == res["checkpoint"][4], "seed stream position changed"
    assert torch.allclose(res["plain"][0], res["checkpoint"][0], atol=1e-6)
    assert torch.allclose(res["plain"][1], res["checkpoint"][1], atol=1e-5), "input gradient changed under recomputation"
    for a, b in zip(res["plain"][2], res["checkpoint"][2]):
        assert torch.allclose(a, b, atol=1e-5), "parameter gradient changed under recomputation"
