#!/bin/bash
# Starts a training script on every node of a hostfile (counterpart of the reference's
# scripts/launch-multinode.sh:1-5, which runs `deepspeed --hostfile ... zero_dp_training.py
# --stage=2 --model_name=facebook/opt-125m` on worker-1).  The first host is the rendezvous address.
# usage: scripts/launch-multinode.sh HOSTFILE [REMOTE_DIR] [-- script args...]
set -euo pipefail
hostfile=${1:?usage: launch-multinode.sh HOSTFILE [REMOTE_DIR] [-- script args...]}
shift
remote_dir="~/distributed-training-b200"
if [ $# -gt 0 ] && [ "$1" != "--" ]; then remote_dir=$1; shift; fi
[ "${1:-}" = "--" ] && shift
[ $# -gt 0 ] || set -- zero_dp_training.py --stage=2 --model-name=facebook/opt-125m
first=$(sed 's/#.*//' "$hostfile" | awk 'NF {print $1; exit}')
scp -q "$hostfile" "$first:$remote_dir/hostfile"
# across nodes the engines use NCCL (peer-mapped memory and NVLS end at the NVSwitch domain)
ssh "$first" "cd $remote_dir && python3 -m dtds_b200.launch --hostfile ./hostfile --ssh --master-port 29500 $*"
