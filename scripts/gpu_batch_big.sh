#!/bin/bash
# The multi-GPU measurement batch for N = 4 or 8 GPUs: all-reduce sweep, fused paths, DDP bench, ZeRO 2/3,
# the multi-billion-parameter ZeRO-3 run, MP / GPipe (BERT base + large) and the library stand-ins.
# Every step has its own (short) timeout and log under gpurun_out/.
N=${1:-8}
BIG=${2:-scale:4096x32}
mkdir -p gpurun_out
run() { name=$1; shift; echo "== $name"; timeout ${TMO:-120} "$@" > gpurun_out/$name.log 2>&1; echo "rc=$? $(grep -v 'frame #\|^\[W\|^W1\|^Exception\|^$\|OMP_NUM\|^\*\*\*\|Warning\|warn\|it/s\]' gpurun_out/$name.log | tail -${TAILN:-1} | cut -c1-${CUTN:-300})"; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
TAILN=12 CUTN=220 run sweep_N$N python pytorch_allreduce.py --world-size $N --sweep --out gpurun_out/allreduce_sweep_${N}gpu.json
run fused_paths_N$N $TR --master-port 29531 bench/fused_paths.py; cp gpurun_out/fused_paths.json gpurun_out/fused_paths_${N}gpu.json 2>/dev/null
CUTN=2600 run bench_N$N $TR --master-port 29533 bench.py --gpus $N --steps 20 --warmup 5
for st in 2 3; do run zero${st}_N$N $TR --master-port 2955$st zero_dp_training.py --stage $st --training-steps $((20 * N)) --quiet; grep -E "Steady|resident" gpurun_out/zero${st}_N$N.log | cut -c1-220; done
TMO=200 run zero3_big_N$N $TR --master-port 29557 zero_dp_training.py --stage 3 --model-name $BIG --training-steps $((8 * N)) --quiet; grep -E "Steady|resident" gpurun_out/zero3_big_N$N.log | cut -c1-240
for model in large base; do
  run pp_graph_${model}_N$N python model_parallel_training.py --device-count $N --model $model --training-steps 30 --pipeline --micro-batch-count 4; grep -E "Steady" gpurun_out/pp_graph_${model}_N$N.log | cut -c1-120
  run mp_naive_${model}_N$N python model_parallel_training.py --device-count $N --model $model --training-steps 30; grep -E "Steady" gpurun_out/mp_naive_${model}_N$N.log | cut -c1-120
done
CUTN=500 run standin_zero3_N$N $TR --master-port 29543 bench/baseline_standins.py zero --stage 3
CUTN=500 run standin_zero2_N$N $TR --master-port 29542 bench/baseline_standins.py zero --stage 2
CUTN=500 run standin_pipe_large_N$N $TR --master-port 29547 bench/baseline_standins.py pipe --model large
CUTN=500 run standin_mp_large_N$N python bench/baseline_standins.py mp --devices $N --model large
