#!/usr/bin/env bash
# Launcher for examples/unet/train.py (capability of the reference's interactive unet/run.sh).  Values come from the
# environment when set, otherwise from a prompt with a default.  Defaults follow the trainer's argparse defaults
# (batch 16, lr 1e-4), not the reference launcher's diverging ones.
set -euo pipefail
validate_ip() { [[ $1 =~ ^([0-9]{1,3}\.){3}[0-9]{1,3}$ ]] || return 1; IFS=. read -r a b c d <<<"$1"; for o in $a $b $c $d; do ((o <= 255)) || return 1; done; }
ask() { local var=$1 prompt=$2 def=$3; if [ -z "${!var:-}" ]; then read -r -p "$prompt [$def]: " v || true; printf -v "$var" '%s' "${v:-$def}"; fi; }
HOST_IP=$(hostname -I 2>/dev/null | awk '{print $1}'); HOST_IP=${HOST_IP:-127.0.0.1}
ask NPROC_PER_NODE "Processes (GPUs) per node" 1
ask NNODES "Number of nodes" 1
ask NODE_RANK "Rank of this node" 0
if [ "$NODE_RANK" = "0" ]; then MASTER_ADDR=${MASTER_ADDR:-$HOST_IP}; else ask MASTER_ADDR "Master node address" ""; fi
validate_ip "$MASTER_ADDR" || { echo "Invalid IP address: $MASTER_ADDR" >&2; exit 1; }
ask MASTER_PORT "Master port" 29500
ask NUM_EPOCHS "Epochs" 100
ask BATCH_SIZE "Batch size per process" 16
ask LEARNING_RATE "Learning rate" 0.0001
ask MODEL_DIR "Model directory" saved_models
ask MODEL_FILENAME "Model filename" model.pth
ask RESUME "Resume from checkpoint? (y/n)" n
ask SYNTHETIC "Synthetic data? (y/n)" n
EXTRA=(); [[ $RESUME =~ ^[Yy] ]] && EXTRA+=(--resume); [[ $SYNTHETIC =~ ^[Yy] ]] && EXTRA+=(--synthetic)
for d in logs "$MODEL_DIR"; do [ -d "$d" ] || { echo "Directory '$d' does not exist. Please create it first." >&2; exit 1; }; done
[[ $SYNTHETIC =~ ^[Yy] ]] || [ -d data ] || { echo "Directory 'data' does not exist." >&2; exit 1; }
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
exec torchrun --nproc_per_node="$NPROC_PER_NODE" --nnodes="$NNODES" --node_rank="$NODE_RANK" \
  --master_addr="$MASTER_ADDR" --master_port="$MASTER_PORT" "$HERE/train.py" \
  --num_epochs "$NUM_EPOCHS" --batch_size "$BATCH_SIZE" --learning_rate "$LEARNING_RATE" \
  --model_dir "$MODEL_DIR" --model_filename "$MODEL_FILENAME" "${EXTRA[@]}"
