#!/bin/bash
# One multi-GPU validation batch (N = number of GPUs, default 2); every step has its own timeout and log.
N=${1:-2}
mkdir -p gpurun_out
run() { name=$1; shift; echo "== $name"; timeout 240 "$@" > gpurun_out/$name.log 2>&1; echo "rc=$? $(grep -v 'frame #\|^\[W\|^W1\|^Exception\|^$\|OMP_NUM\|^\*\*\*' gpurun_out/$name.log | tail -${TAILN:-3} | cut -c1-${CUTN:-600})"; }
W="python -u scripts/run_worker.py"
run ddp_fused_graph $W ddp_training_loop $N nccl 200 1 1
run zero_stages_bloom $W zero_stages_match_stage0 $N nccl 220 bloom-tiny
run zero_stages_bert $W zero_stages_match_stage0 $N nccl 220 bert-tiny
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
CUTN=2500 run bench_N$N $TR --master-port 29533 bench.py --gpus $N --steps 20 --warmup 5
TAILN=16 CUTN=160 run timeline_N$N $TR --master-port 29534 bench/step_timeline.py --gpus $N
cp gpurun_out/step_timeline_N$N.txt gpurun_out/step_timeline_dense_N$N.txt 2>/dev/null
TAILN=6 run zero2_graph $TR --master-port 29537 zero_dp_training.py --stage 2 --training-steps 24 --quiet
DTDS_ZERO3_GRAPH=0 TAILN=6 run zero3_eager $TR --master-port 29538 zero_dp_training.py --stage 3 --training-steps 24 --quiet
TAILN=6 run zero3_graph $TR --master-port 29539 zero_dp_training.py --stage 3 --training-steps 24 --quiet
