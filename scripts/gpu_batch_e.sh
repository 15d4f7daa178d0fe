#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
run() { name=$1; shift; echo "== $name"; timeout ${TMO:-200} "$@" > gpurun_out/$name.log 2>&1; echo "rc=$? $(grep -v 'frame #\|^\[W\|^W1\|^Exception\|^$\|OMP_NUM\|^\*\*\*\|Warning\|warn\|it/s\]' gpurun_out/$name.log | tail -${TAILN:-2} | cut -c1-${CUTN:-400})"; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
TMO=150 run zero_stages_bloom python -u scripts/run_worker.py zero_stages_match_stage0 $N nccl 140 bloom-tiny
TMO=300 TAILN=3 run zero3_big_N$N $TR --master-port 29557 zero_dp_training.py --stage 3 --model-name scale:4096x32 --training-steps $((8 * N)) --quiet; grep -E "Steady|resident|parameter count" gpurun_out/zero3_big_N$N.log | cut -c1-260
CUTN=700 TAILN=1 run bench_N$N $TR --master-port 29533 bench.py --gpus $N --steps 20 --warmup 5
TAILN=14 CUTN=160 run timeline_N$N $TR --master-port 29534 bench/step_timeline.py --gpus $N
TAILN=3 run zero3_N$N $TR --master-port 29553 zero_dp_training.py --stage 3 --training-steps $((12 * N)) --quiet; grep -E "Steady|resident" gpurun_out/zero3_N$N.log | cut -c1-220
TMO=150 run agemm python -u scripts/run_worker.py allgather_gemm_matches_gathered $N nccl 120
