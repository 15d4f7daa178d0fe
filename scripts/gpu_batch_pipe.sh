#!/bin/bash
# Pipeline / model-parallel validation + timings on N GPUs (default 2).
N=${1:-2}
mkdir -p gpurun_out
run() { name=$1; shift; echo "== $name"; timeout 300 "$@" > gpurun_out/$name.log 2>&1; echo "rc=$? $(grep -v 'frame #\|^\[W\|^W1\|^Exception\|^$\|OMP_NUM\|^\*\*\*\|Warning' gpurun_out/$name.log | tail -${TAILN:-3} | cut -c1-${CUTN:-400})"; }
W="python -u scripts/run_worker.py"
run pipe_equiv_graphs $W pipeline_matches_single_process $N nccl 200 4
run pipe_recompute $W pipeline_recompute_replays_dropout $N nccl 200 4
run pipe_resume $W pipeline_checkpoint_resume $N nccl 200 2
TR="python"
for model in base large; do
  TAILN=14 run mp_naive_${model}_N$N $TR model_parallel_training.py --device-count $N --model $model --training-steps 30
  TAILN=14 run pp_graph_${model}_N$N $TR model_parallel_training.py --device-count $N --model $model --training-steps 30 --pipeline --micro-batch-count 4
  DTDS_PIPE_GRAPHS=0 TAILN=6 run pp_eager_${model}_N$N $TR model_parallel_training.py --device-count $N --model $model --training-steps 30 --pipeline --micro-batch-count 4
done
