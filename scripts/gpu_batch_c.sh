#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
run() { name=$1; shift; echo "== $name"; timeout 300 "$@" > gpurun_out/$name.log 2>&1; echo "rc=$? $(grep -v 'frame #\|^\[W\|^W1\|^Exception\|^$\|OMP_NUM\|^\*\*\*\|Warning\|warn' gpurun_out/$name.log | tail -${TAILN:-2} | cut -c1-${CUTN:-400})"; }
W="python -u scripts/run_worker.py"
run pipe_graphs_never $W pipeline_graphs_match_eager $N nccl 250 4 never
run pipe_graphs_except_last $W pipeline_graphs_match_eager $N nccl 250 4 except_last
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
CUTN=2600 TAILN=1 run bench_N$N $TR --master-port 29533 bench.py --gpus $N --steps 20 --warmup 5
TAILN=9 run pp_graph_base_N$N python model_parallel_training.py --device-count $N --training-steps 30 --pipeline --micro-batch-count 4
for st in 2 3; do TAILN=1 CUTN=600 run standin_zero${st}_N$N $TR --master-port 2954$st bench/baseline_standins.py zero --stage $st; done
TAILN=1 CUTN=600 run standin_mp_N$N python bench/baseline_standins.py mp --devices $N
TAILN=1 CUTN=600 run standin_pipe_N$N $TR --master-port 29547 bench/baseline_standins.py pipe
