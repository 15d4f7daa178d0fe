#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
run() { name=$1; shift; echo "== $name"; timeout ${TMO:-200} "$@" > gpurun_out/$name.log 2>&1; echo "rc=$? $(grep -v 'frame #\|^\[W\|^W1\|^Exception\|^$\|OMP_NUM\|^\*\*\*\|Warning\|warn\|it/s\]' gpurun_out/$name.log | tail -${TAILN:-2} | cut -c1-${CUTN:-400})"; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
TMO=150 run pipe_graphs_never python -u scripts/run_worker.py pipeline_graphs_match_eager $N nccl 120 4 never
TMO=150 run pipe_graphs_except_last python -u scripts/run_worker.py pipeline_graphs_match_eager $N nccl 120 4 except_last
TMO=300 TAILN=3 run zero3_big_N$N $TR --master-port 29557 zero_dp_training.py --stage 3 --model-name scale:4096x32 --training-steps $((8 * N)) --quiet; grep -E "Steady|resident|parameter count" gpurun_out/zero3_big_N$N.log | cut -c1-260
TMO=300 CUTN=1500 TAILN=1 run ref_arm_N$N $TR --master-port 29561 bench.py --impl reference --gpus $N --steps 20 --warmup 5
