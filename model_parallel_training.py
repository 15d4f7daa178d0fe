"""Layer-split model-parallel / GPipe pipeline-parallel BERT training.

Command line of the reference: ``--pipeline``, ``--verbose``, ``--batch-size`` (global),
``--training-steps``, ``--device-count``, ``--micro-batch-count``; output: total training time and
the per-device average idle time table.  On GPUs every device is driven by its own process and
neighbouring stages exchange activations and gradients with NCCL send/recv overlapped with
micro-batch compute; with one device (or on CPU with ``--device-count 1``) everything runs in this
process exactly in the reference's order: logits -> CrossEntropyLoss -> backward -> step.

    python model_parallel_training.py --device-count 4                       # naive model parallel
    python model_parallel_training.py --device-count 4 --pipeline --micro-batch-count 4
    python model_parallel_training.py --backend gloo --device-count 2 --model tiny --pipeline
"""
import argparse
import os
import time

import torch
import torch.multiprocessing as mp
from tabulate import tabulate
from torch.utils.data import DataLoader
from tqdm import tqdm

from model.bert_mp import BertModelWithMP
from util import get_device_count, load_wikitext, synthetic_collator, synthetic_tokenizer


def summarize_idle_time(bert, training_steps):
    rows = [["Device", "Average Idle Time (ms)"]]
    for device, (total_ms, _count) in sorted(bert.collect_idle_time().items()):
        rows.append([device, total_ms / training_steps])
    return tabulate(rows, headers="firstrow", floatfmt=".2f", tablefmt="fancy_grid")


def run(rank, device_count, args, backend):
    from dtds_b200 import comm
    from dtds_b200.models import get_config
    from dtds_b200.ops import functional as Fn
    from dtds_b200.optim import AdamW

    distributed = device_count > 1 and backend is not None
    if distributed:
        if backend == "nccl":
            torch.cuda.set_device(rank)
        comm.init_process_group(backend=backend, rank=rank, world_size=device_count,
                                master_addr="127.0.0.1", master_port=args.port)
    config = get_config({"base": "bert-base-cased", "large": "bert-large-cased", "tiny": "bert-tiny"}[args.model])
    torch.manual_seed(0)
    bert = BertModelWithMP(config=config, device_count=device_count, verbose=args.verbose)
    model = bert.to_pipeline(chunks=args.micro_batch_count, checkpoint=args.checkpoint) if args.pipeline else bert

    tokenizer = synthetic_tokenizer(config)
    collator = synthetic_collator(tokenizer, mlm=True)
    params = [p for p in bert.parameters()]
    optimizer = AdamW(params, lr=5e-5, weight_decay=0.01) if params else None

    seq = args.seq_len or tokenizer.model_max_length
    train_dataset = load_wikitext(tokenizer, collator, max_length=seq, path=args.data)
    generator = torch.Generator().manual_seed(1)   # same shuffle on every stage
    train_dataloader = DataLoader(train_dataset.select(range(args.batch_size * args.training_steps)),
                                  shuffle=True, batch_size=args.batch_size, generator=generator)
    bert.train()
    last = (not distributed) or rank == device_count - 1
    training_start_time = time.time()
    progress_bar = tqdm(range(args.training_steps)) if (not args.verbose and last) else None
    steady_start, steady_from = None, min(5, max(args.training_steps - 1, 0))   # eager warm-up + graph capture are over
    for step_idx, batch in enumerate(train_dataloader):
        if step_idx == steady_from:   # connection set-up, autotuning and allocator growth are over
            if torch.cuda.is_available():
                torch.cuda.synchronize()
            steady_start = time.time()
            for tr in bert.trackers.values():
                tr.reset()
        if distributed:
            bert.forward_backward(batch["input_ids"], batch["labels"])
        else:
            input_ids = batch["input_ids"].to(bert.embedding_device)
            outputs = model(input_ids)
            if args.pipeline:
                outputs = outputs.local_value()
            labels = batch["labels"].to(bert.head_device)
            loss = Fn.cross_entropy(outputs, labels)
            loss.backward()
        if optimizer is not None:
            optimizer.step()
            optimizer.zero_grad()
        if progress_bar is not None:
            progress_bar.update(1)
    if torch.cuda.is_available():
        torch.cuda.synchronize()
    training_end_time = time.time()
    steady_steps = max(args.training_steps - steady_from, 1)
    table = summarize_idle_time(bert, steady_steps)
    if last:
        print(f"\nTotal Training Time: {training_end_time - training_start_time:.2f} seconds")
        if steady_start is not None:
            dt = training_end_time - steady_start
            print(f"Steady state (steps {steady_from + 1}..{args.training_steps}): {dt / steady_steps * 1e3:.2f} ms/step, "
                  f"{args.batch_size * steady_steps / dt:.1f} samples/s; idle times below exclude the first {steady_from} steps")
        print("\nAverage Idle Time per Device:")
        print(table)
    if distributed:
        comm.destroy_process_group()


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    parser.add_argument("--pipeline", action="store_true")
    parser.add_argument("--verbose", action="store_true")
    parser.add_argument("--batch-size", type=int, default=16)
    parser.add_argument("--training-steps", type=int, default=250)
    parser.add_argument("--device-count", type=int, default=None)
    parser.add_argument("--micro-batch-count", type=int, default=4)
    parser.add_argument("--checkpoint", choices=["never", "except_last", "always"], default="never",
                        help="activation recomputation in --pipeline mode (torch Pipe's default was except_last)")
    # additions (not in the reference)
    parser.add_argument("--model", choices=["base", "large", "tiny"], default="base")
    parser.add_argument("--backend", choices=["nccl", "gloo"], default=None)
    parser.add_argument("--port", type=int, default=12356)
    parser.add_argument("--seq-len", type=int, default=None)
    parser.add_argument("--data", type=str, default=None,
                        help="local text file, one passage per line (default: synthetic wikitext-shaped rows)")
    args = parser.parse_args()

    backend = args.backend or ("nccl" if torch.cuda.is_available() else None)
    device_count = args.device_count or get_device_count()
    if backend == "nccl" and device_count > get_device_count():
        print(f"Cannot use {device_count} cuda devices. Only {get_device_count()} devices are available. "
              f"Training will be done using {get_device_count()} devices.")
        device_count = get_device_count()
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if device_count > 1 and backend is not None:
        mp.spawn(run, args=(device_count, args, backend), nprocs=device_count, join=True)
    else:
        run(0, 1, args, None)
