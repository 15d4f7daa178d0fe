"""ZeRO data-parallel causal-LM training (stages 0-3) behind the DeepSpeed-shaped engine.

Command line of the reference (``--model-name``, ``--batch-size`` per GPU, ``--training-steps``,
``--stage``; unknown arguments such as the launcher's ``--local_rank`` are swallowed) and the same
console output: the stage line, the optimizer line, a MEMSTATS line per step on rank 0, the total
training time, the per-collective summary table and the total communication latency.

Launch one process per GPU with the in-repo launcher (or torchrun):

    python -m dtds_b200.launch --nproc-per-node 8 zero_dp_training.py --stage 3
    python zero_dp_training.py --stage 0                         # single process

``--model-name`` takes the in-repo configurations (bigscience/bloom-560m, facebook/opt-125m,
bloom-tiny, ...) or ``scale:<hidden>x<layers>`` for a hand-sized multi-billion-parameter decoder.
Weights are random-init and the rows synthetic (no hub access).  Unlike the reference the
launcher's ``--model_name`` spelling is accepted too (its scripts/launch-multinode.sh passes the
underscore form, which argparse silently dropped: quirk Q3), and latencies are labelled in the
unit they are measured in (milliseconds; quirk Q2).
"""
import argparse
import os
import time

import torch
from torch.utils.data import DataLoader
from torch.utils.data.distributed import DistributedSampler
from tqdm import tqdm

from util import load_wikitext, synthetic_collator, synthetic_tokenizer


def resolve_config(model_name):
    from dtds_b200.models import get_config, transformer_scale_config
    if model_name.startswith("scale:"):
        hidden, layers = model_name.split(":", 1)[1].lower().split("x")
        return transformer_scale_config(int(hidden), int(layers))
    return get_config(model_name)


def train(model_name, batch_size, training_steps, stage, *args, bf16=None, seq_len=512, quiet=False,
          cuda_graph=False, pretrained=None, data=None, save_dir=None, resume=None):
    from dtds_b200 import comm
    from dtds_b200.models import build_model
    from dtds_b200.parallel import zero
    from dtds_b200.utils import memory_status

    rank = int(os.getenv("LOCAL_RANK", "0"))
    world_size = int(os.getenv("WORLD_SIZE", "1"))
    on_gpu = torch.cuda.is_available()
    device = torch.device("cuda", rank) if on_gpu else torch.device("cpu")
    if world_size > 1:
        comm.init_process_group()

    cfg = resolve_config(model_name)
    torch.manual_seed(0)
    if model_name.startswith("scale:") and on_gpu and bf16 is not False:
        # multi-billion-parameter shapes are created on the GPU in bf16 directly: an fp32 host copy
        # per rank (the reference's from_pretrained on the CPU) would be tens of GB times the ranks
        old = torch.get_default_dtype()
        torch.set_default_dtype(torch.bfloat16)
        try:
            with torch.device(device):
                model = build_model(cfg)
        finally:
            torch.set_default_dtype(old)
    else:
        model = build_model(cfg)
    if pretrained:   # local HF checkpoint of the same architecture (the reference downloads it)
        from dtds_b200.models import load_pretrained
        load_pretrained(model, pretrained)

    ds_config = {
        "train_micro_batch_size_per_gpu": batch_size,
        "optimizer": {"type": "Adam", "params": {"lr": 0.00015}},
        "comms_logger": {"enabled": True, "verbose": False, "prof_all": True, "debug": False},
        "zero_optimization": {"stage": stage, "reduce_bucket_size": 5e6},
        "bf16": {"enabled": on_gpu if bf16 is None else bf16},
    }
    model_engine, optimizer, _, _ = zero.initialize(model=model, model_parameters=model.parameters(),
                                                    config=ds_config)
    print(f"Device {rank} - ZeRO Stage: {model_engine.zero_optimization_stage()}")
    group = optimizer.param_groups[0]
    print(f"Device {rank} - Optimizer: lr={group['lr']}; betas={group['betas']}; eps={group['eps']}; "
          f"parameter count={sum(p._numel for p in group['params']):,}")

    if rank == 0:
        rep = model_engine.memory_report()
        gb = 1024 ** 3
        print(f"Device {rank} - ZeRO-{stage} resident model state per rank ({world_size} rank(s)): "
              f"parameters {rep['parameters'] / gb:.3f} GB, gradients {rep['gradients'] / gb:.3f} GB, "
              f"optimizer {rep['optimizer'] / gb:.3f} GB; device memory in use {rep['device_used'] / gb:.2f} GB")
    tokenizer = synthetic_tokenizer(cfg)
    collator = synthetic_collator(tokenizer, mlm=False)
    seq = min(seq_len, cfg.max_position_embeddings)
    train_dataset = load_wikitext(tokenizer, collator, max_length=seq, path=data).select(
        range(batch_size * training_steps))
    train_dataloader = DataLoader(train_dataset, batch_size=batch_size, shuffle=False,
                                  sampler=DistributedSampler(train_dataset, num_replicas=world_size,
                                                             rank=comm.get_rank()))
    model_engine.train()
    comm.comms_logger.resolve()
    comm.comms_logger.comms_dict = {}     # only the training loop counts towards the totals below

    training_start_time = time.time()
    progress_bar = tqdm(range(training_steps)) if rank == 0 and not quiet else None
    def step_fn(input_ids, labels):
        loss = model_engine(input_ids, labels=labels).loss
        model_engine.backward(loss)
        model_engine.step()
        return loss

    from dtds_b200.utils.trace import nvtx_range
    start_step = 0
    if resume:   # model.pt (consolidated fp32) + this rank's optimizer partition + step counter
        from dtds_b200.utils.ckpt import load_checkpoint
        start_step = load_checkpoint(resume, model_engine, optimizer)
        if rank == 0:
            print(f"Resumed from {resume} at step {start_step}")
    graphed = None
    # The whole step (forward, backward, side-stream bucket kernels, optimizer) is replayed as one CUDA
    # graph.  Collectives inside a replay cannot be timed one by one: the comm table below then
    # holds the eager warm-up steps before the capture.  Stage 3 is captured too: the per-block
    # parameter gathers the module hooks issue land in the graph with their (pool) addresses
    # (DTDS_ZERO3_GRAPH=0 keeps stage 3 eager).
    use_graph = cuda_graph and on_gpu and (stage < 3 or os.environ.get("DTDS_ZERO3_GRAPH", "1") != "0")
    steady_start, steady_from = None, 3
    for step_idx, batch in enumerate(train_dataloader):
        if step_idx < start_step:
            continue   # rows the interrupted run already consumed
        if step_idx == steady_from:
            if on_gpu:
                torch.cuda.synchronize()
            steady_start = time.time()
        input_ids = batch["input_ids"].to(device)
        labels = batch["labels"].to(device)
        with nvtx_range(f"train step {step_idx}"):
            if use_graph and input_ids.shape[0] == batch_size:
                if graphed is None:   # capture forward + backward + fused ZeRO step kernels once
                    from dtds_b200.parallel.graph import GraphedStep
                    graphed = GraphedStep(step_fn, [input_ids, labels], warmup=2)
                loss = graphed(input_ids, labels)
            else:
                loss = step_fn(input_ids, labels)
        if rank == 0:
            if not quiet:
                memory_status("Memory stats after training step:")
            if progress_bar is not None:
                progress_bar.update(1)
    if on_gpu:
        torch.cuda.synchronize()
    training_end_time = time.time()
    if save_dir:
        from dtds_b200.utils.ckpt import save_checkpoint
        save_checkpoint(save_dir, model_engine, optimizer, step=len(train_dataloader))
        if rank == 0:
            print(f"Saved checkpoint to {save_dir}")
    if rank == 0:
        print(f"\nTotal Training Time: {training_end_time - training_start_time:.2f} seconds")
        n_local = len(train_dataloader)
        if steady_start is not None and n_local > steady_from:
            dt = (training_end_time - steady_start) / (n_local - steady_from)
            print(f"Steady state (local steps {steady_from + 1}..{n_local}): {dt * 1e3:.2f} ms/step, "
                  f"{batch_size * world_size / dt:.1f} samples/s over {world_size} rank(s)")

    total_comms_latency = comm.comms_logger.total_latency_ms()
    comm.log_summary()
    if rank == 0:
        print(f"\nTotal Communication Latency: {total_comms_latency:.2f} ms "
              f"({total_comms_latency / 1e3:.3f} seconds)")
    comm.destroy_process_group()
    return model_engine


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    parser.add_argument("--model-name", "--model_name", dest="model_name", type=str, default="bigscience/bloom-560m")
    parser.add_argument("--batch-size", type=int, default=1)
    parser.add_argument("--training-steps", type=int, default=100)
    parser.add_argument("--stage", type=int, default=0)
    parser.add_argument("--seq-len", type=int, default=512)
    parser.add_argument("--fp32", action="store_true", help="keep fp32 parameters (library collectives)")
    parser.add_argument("--quiet", action="store_true", help="no per-step MEMSTATS / progress bar")
    parser.add_argument("--cuda-graph", dest="cuda_graph", action="store_true", default=True,
                        help="replay the step as one CUDA graph (default on a GPU)")
    parser.add_argument("--no-cuda-graph", dest="cuda_graph", action="store_false",
                        help="launch every kernel eagerly (every collective is then timed in the comm table)")
    parser.add_argument("--save-dir", type=str, default=None, help="write a checkpoint here after the last step")
    parser.add_argument("--resume", type=str, default=None, help="checkpoint directory to continue from")
    parser.add_argument("--data", type=str, default=None,
                        help="local text file, one passage per line (default: synthetic wikitext-shaped rows)")
    parser.add_argument("--pretrained", type=str, default=None,
                        help="local HF checkpoint (file or directory) to start from instead of random weights")
    args, extra_args = parser.parse_known_args()
    train(args.model_name, args.batch_size, args.training_steps, args.stage, *extra_args,
          bf16=False if args.fp32 else None, seq_len=args.seq_len, quiet=args.quiet, cuda_graph=args.cuda_graph,
          pretrained=args.pretrained, data=args.data, save_dir=args.save_dir, resume=args.resume)
