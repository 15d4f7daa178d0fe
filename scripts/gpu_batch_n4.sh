#!/bin/bash
N=${1:-4}
mkdir -p gpurun_out
run() { name=$1; shift; echo "== $name"; timeout ${TMO:-120} "$@" > gpurun_out/$name.log 2>&1; echo "rc=$? $(grep -v 'frame #\|^\[W\|^W1\|^Exception\|^$\|OMP_NUM\|^\*\*\*\|Warning\|warn\|it/s\]' gpurun_out/$name.log | tail -${TAILN:-1} | cut -c1-${CUTN:-300})"; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
TAILN=12 CUTN=220 run sweep_N$N python pytorch_allreduce.py --world-size $N --sweep --out gpurun_out/allreduce_sweep_${N}gpu.json
CUTN=900 run bench_N$N $TR --master-port 29533 bench.py --gpus $N --steps 20 --warmup 5
DTDS_NVLS=0 CUTN=400 run bench_p2p_N$N $TR --master-port 29535 bench.py --gpus $N --steps 20 --warmup 5
TAILN=14 CUTN=160 run timeline_N$N $TR --master-port 29534 bench/step_timeline.py --gpus $N
run fused_paths_N$N $TR --master-port 29531 bench/fused_paths.py; cp gpurun_out/fused_paths.json gpurun_out/fused_paths_${N}gpu.json 2>/dev/null
run pp_graph_large_N$N python model_parallel_training.py --device-count $N --model large --training-steps 30 --pipeline --micro-batch-count 4; grep -E "Steady" gpurun_out/pp_graph_large_N$N.log | cut -c1-120
run mp_naive_large_N$N python model_parallel_training.py --device-count $N --model large --training-steps 30; grep -E "Steady" gpurun_out/mp_naive_large_N$N.log | cut -c1-120
run pp_graph_base_N$N python model_parallel_training.py --device-count $N --model base --training-steps 30 --pipeline --micro-batch-count 4; grep -E "Steady" gpurun_out/pp_graph_base_N$N.log | cut -c1-120
