"""Data-parallel BERT masked-LM training (DDP).

Same command line as the reference script (``--batch-size`` per rank, ``--training-steps``,
``--device-count``, ``--bucket-size`` MiB, ``--model base|large``) and the same final line
``Total Training Time: X seconds`` from every rank.  What runs underneath is this repository's
engine: the in-repo BERT built on sm_100a kernels (random init — no hub access), synthetic
wikitext-shaped MLM rows sharded by ``DistributedSampler``, gradients in a flat symmetric-memory
buffer whose buckets are all-reduced by P2P kernels while backward is still running, a one-launch
fused AdamW (lr 5e-5), and optionally the whole step replayed as a CUDA graph.

    python data_parallel_training.py --model base --batch-size 4 --training-steps 1000
    python data_parallel_training.py --backend gloo --device-count 2 --model tiny    # CPU plumbing
    python data_parallel_training.py --model block --backend gloo --device-count 2   # TransformerBlock
"""
import argparse
import os
import time

import torch
import torch.multiprocessing as mp
from torch.utils.data import DataLoader
from torch.utils.data.distributed import DistributedSampler
from tqdm import tqdm

from util import get_device_count, load_wikitext, synthetic_collator, synthetic_tokenizer

MODEL_NAMES = {"base": "bert-base-cased", "large": "bert-large-cased", "tiny": "bert-tiny"}


def create_process_group(rank, world_size, backend, port):
    from dtds_b200 import comm
    comm.init_process_group(backend=backend, rank=rank, world_size=world_size,
                            master_addr="127.0.0.1", master_port=port)


class _BlockRegressor(torch.nn.Module):
    """``--model block``: the hookable TransformerBlock (model/transformer.py) with a toy
    embedding and an MSE objective — the CPU/gloo plumbing configuration of BASELINE.json."""

    def __init__(self, hidden=64, heads=4, vocab=1000):
        super().__init__()
        from types import SimpleNamespace
        from model.transformer import TransformerBlock
        cfg = SimpleNamespace(hidden_size=hidden, num_attention_heads=heads, ffn_dim=4 * hidden,
                              enable_bias=True, dropout=0.0, max_position_embeddings=512)
        self.embed = torch.nn.Embedding(vocab, hidden)
        self.block = TransformerBlock(cfg)
        self.out = torch.nn.Linear(hidden, vocab)

    def forward(self, input_ids, labels=None):
        from types import SimpleNamespace
        logits = self.out(self.block(self.embed(input_ids)))
        loss = None
        if labels is not None:
            loss = torch.nn.functional.cross_entropy(logits.view(-1, logits.shape[-1]).float(), labels.view(-1),
                                                     ignore_index=-100)
        return SimpleNamespace(loss=loss, logits=logits)


def train(rank, world_size, batch_size, training_steps, bucket_size, model, backend, port, use_graph, seq_len,
          overlap_optimizer=False, pretrained=None, data=None, save_dir=None, resume=None, json_out=None):
    from dtds_b200.models import build_model, get_config
    from dtds_b200.optim import HFAdamW
    from dtds_b200.parallel.ddp import DistributedDataParallel

    on_gpu = backend == "nccl"
    if on_gpu:
        torch.cuda.set_device(rank)
    create_process_group(rank, world_size, backend, port)
    device = torch.device("cuda", rank) if on_gpu else torch.device("cpu")

    torch.manual_seed(0)
    if model == "block":
        cfg = get_config("bert-tiny")
        net = _BlockRegressor(vocab=cfg.vocab_size).to(device)
    else:
        cfg = get_config(MODEL_NAMES[model])
        net = build_model(cfg)
        if pretrained:   # local HF BertForMaskedLM checkpoint (the reference downloads bert-*-cased)
            from dtds_b200.models import load_pretrained
            load_pretrained(net, pretrained)
        net = net.to(device=device, dtype=torch.bfloat16 if on_gpu else torch.float32)
    tokenizer = synthetic_tokenizer(cfg)
    collator = synthetic_collator(tokenizer, mlm=True)
    optimizer = HFAdamW(net.parameters(), lr=5e-5)

    train_dataset = load_wikitext(tokenizer, collator, max_length=seq_len, path=data).select(
        range(batch_size * training_steps))
    train_dataloader = DataLoader(train_dataset, batch_size=batch_size, shuffle=False,
                                  sampler=DistributedSampler(train_dataset), pin_memory=on_gpu)

    ddp = DistributedDataParallel(net, device_ids=[rank] if on_gpu else None, bucket_cap_mb=bucket_size)
    ddp.train()
    if overlap_optimizer:
        ddp.fuse_optimizer(optimizer)   # Adam per bucket right behind its all-reduce, under backward

    def step_fn(input_ids, labels):
        optimizer.zero_grad()
        loss = ddp(input_ids, labels=labels).loss
        loss.backward()
        optimizer.step()
        return loss

    start_step = 0
    if resume:
        from dtds_b200.utils.ckpt import load_checkpoint
        start_step = load_checkpoint(resume, ddp, optimizer)
        if rank == 0:
            print(f"Resumed from {resume} at step {start_step}")
    graphed = None
    training_start_time = time.time()
    progress_bar = tqdm(range(training_steps), disable=rank != 0)
    from dtds_b200.utils.trace import annotate_model, nvtx_enabled, nvtx_range
    if nvtx_enabled():   # DTDS_NVTX=1: per-step / per-layer / per-bucket ranges for a profiler timeline
        annotate_model(net)
    steady_start, steady_from = None, start_step + 3   # set-up, autotuning, graph capture are over by then
    for step_idx, batch in enumerate(train_dataloader):
        if step_idx < start_step:
            continue   # rows the interrupted run already consumed
        if step_idx == steady_from:
            if on_gpu:
                torch.cuda.synchronize()
            steady_start = time.time()
        input_ids, labels = batch["input_ids"], batch["labels"]
        with nvtx_range(f"train step {step_idx}"):
            if use_graph and on_gpu and input_ids.shape[0] == batch_size:
                if graphed is None:
                    from dtds_b200.parallel.graph import GraphedStep
                    graphed = GraphedStep(step_fn, [input_ids.to(device), labels.to(device)], warmup=2)
                graphed(input_ids, labels)
            else:
                step_fn(input_ids.to(device, non_blocking=True), labels.to(device, non_blocking=True))
        progress_bar.update(1)
    if on_gpu:
        torch.cuda.synchronize()
    training_end_time = time.time()
    if save_dir:
        from dtds_b200.utils.ckpt import save_checkpoint
        save_checkpoint(save_dir, ddp, optimizer, step=len(train_dataloader))
    print(f"\nTotal Training Time: {training_end_time - training_start_time:.2f} seconds")
    n_local = len(train_dataloader)
    if rank == 0 and steady_start is not None and n_local > steady_from:
        dt = (training_end_time - steady_start) / (n_local - steady_from)
        line = {"ms_per_step": round(dt * 1e3, 3), "samples_per_s": round(batch_size * world_size / dt, 1),
                "ranks": world_size, "steps_timed": n_local - steady_from}
        print(f"Steady state (local steps {steady_from + 1}..{n_local}): {line['ms_per_step']:.2f} ms/step, "
              f"{line['samples_per_s']:.1f} samples/s over {world_size} rank(s)")
        if json_out:
            import json
            with open(json_out, "w") as f:
                json.dump({"script": "data_parallel_training.py", "model": model, "batch_size": batch_size,
                           "cuda_graph": bool(use_graph), **line}, f)
    from dtds_b200 import comm
    comm.destroy_process_group()


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    parser.add_argument("--batch-size", type=int, default=4)
    parser.add_argument("--training-steps", type=int, default=1000)
    parser.add_argument("--device-count", type=int, default=None)
    parser.add_argument("--bucket-size", type=int, default=25)
    parser.add_argument("--model", type=str, choices=["base", "large", "tiny", "block"], default="base")
    # additions (not in the reference): transport, port, sequence length, CUDA-graph replay
    parser.add_argument("--backend", choices=["nccl", "gloo"], default=None)
    parser.add_argument("--port", type=int, default=12355)
    parser.add_argument("--seq-len", type=int, default=None)
    parser.add_argument("--cuda-graph", action="store_true")
    parser.add_argument("--overlap-optimizer", action="store_true")
    parser.add_argument("--json-out", type=str, default=None, help="write the steady-state throughput here (rank 0)")
    parser.add_argument("--save-dir", type=str, default=None, help="write a checkpoint here after the last step")
    parser.add_argument("--resume", type=str, default=None, help="checkpoint directory to continue from")
    parser.add_argument("--data", type=str, default=None,
                        help="local text file, one passage per line (default: synthetic wikitext-shaped rows)")
    parser.add_argument("--pretrained", type=str, default=None,
                        help="local HF checkpoint (file or directory) to start from instead of random weights")
    args = parser.parse_args()

    backend = args.backend or ("nccl" if torch.cuda.is_available() else "gloo")
    device_count = args.device_count or (get_device_count() if backend == "nccl" else 1)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    mp.spawn(train,
             args=(device_count, args.batch_size, args.training_steps, args.bucket_size, args.model, backend,
                   args.port, args.cuda_graph, args.seq_len, args.overlap_optimizer, args.pretrained, args.data,
                   args.save_dir, args.resume, args.json_out),
             nprocs=device_count, join=True)
