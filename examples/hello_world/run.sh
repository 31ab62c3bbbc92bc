#!/usr/bin/env bash
# Launcher for examples/hello_world/hello_world.py.  Every value can come from the environment (non-interactive) or is
# prompted for with a default.
set -euo pipefail
ask() { local var=$1 prompt=$2 def=$3; if [ -z "${!var:-}" ]; then read -r -p "$prompt [$def]: " v || true; printf -v "$var" '%s' "${v:-$def}"; fi; }
ask NPROC_PER_NODE "Processes (GPUs) per node" 2
ask NNODES "Number of nodes" 1
ask NODE_RANK "Rank of this node" 0
ask MASTER_ADDR "Master address" 127.0.0.1
ask MASTER_PORT "Master port" 29500
ask BACKEND "Backend (nccl|gloo)" nccl
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
exec torchrun --nproc_per_node="$NPROC_PER_NODE" --nnodes="$NNODES" --node_rank="$NODE_RANK" \
  --master_addr="$MASTER_ADDR" --master_port="$MASTER_PORT" "$HERE/hello_world.py" --backend "$BACKEND"
