#!/bin/bash
# 2-GPU bench over DDP bucket sizes / optimizer placement.
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 2 --steps 20 --warmup 5 "${@:2}" 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('${*:2}', d['ms_per_step'], d['e2e']['ms_per_step'])"; }
run 29521 --bucket-size 10
run 29522 --bucket-size 50
run 29523 --no-overlap-optimizer
