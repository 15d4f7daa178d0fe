#!/bin/bash
# Prepares the nodes of a hostfile for `scripts/launch-multinode.sh`:
#   1. one key pair for node-to-node ssh, installed on every node (the launcher starts itself on
#      each node over ssh from the first one);
#   2. this checkout copied to the same path on every node and the CUDA library built there;
#   3. every node's ~/.ssh/config learns the other nodes' names.
# usage: scripts/setup-workers.sh HOSTFILE [REMOTE_DIR]      (hosts must already accept your ssh key)
set -euo pipefail
hostfile=${1:?usage: setup-workers.sh HOSTFILE [REMOTE_DIR]}
remote_dir=${2:-\~/distributed-training-b200}
repo=$(cd "$(dirname "$0")/.." && pwd)
mapfile -t hosts < <(sed 's/#.*//' "$hostfile" | awk 'NF {print $1}')
[ "${#hosts[@]}" -gt 0 ] || { echo "no hosts in $hostfile"; exit 1; }

keys=$(mktemp -d)
ssh-keygen -q -t ed25519 -N "" -f "$keys/id_cluster"

setup_node() {
  local h=$1
  scp -q "$keys/id_cluster" "$keys/id_cluster.pub" "$h:.ssh/"
  ssh "$h" 'cat ~/.ssh/id_cluster.pub >> ~/.ssh/authorized_keys && chmod 600 ~/.ssh/id_cluster'
  rsync -a --delete --exclude .git --exclude gpurun_out --exclude '*.so' --exclude '__pycache__' "$repo/" "$h:$remote_dir/"
  ssh "$h" "cd $remote_dir && scripts/worker-prereqs.sh && python3 -c 'import __graft_entry__ as g; g.build()'"
}
write_config() {
  local h=$1
  { for peer in "${hosts[@]}"; do
      printf 'Host %s\n    IdentityFile ~/.ssh/id_cluster\n    StrictHostKeyChecking no\n\n' "$peer"
    done; } | ssh "$h" 'cat >> ~/.ssh/config'
}
for h in "${hosts[@]}"; do setup_node "$h" & done
wait
for h in "${hosts[@]}"; do write_config "$h"; done
rm -rf "$keys"
echo "ready: ${hosts[*]}"
