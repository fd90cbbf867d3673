#!/usr/bin/env bash
# Multi-node launch: one detached tmux session per host running torch.distributed.run with a c10d
# rendezvous on the first host (same shape as the reference's 05-training-llama-405b/launch.sh).
#   bash launch.sh [hosts-file]        monitor: find ../logs -name '*stderr.log' | xargs tail -f
#   kill:  xargs -a hosts -I{} ssh {} tmux kill-session -t torchrun-llama-405b
set -euo pipefail
HOSTS="${1:-hosts}"
HEAD="$(head -n 1 "$HOSTS")"
NNODES="$(grep -c '^' "$HOSTS")"
SESSION=torchrun-llama-405b
EXPERIMENT="${EXPERIMENT:-llama-405b}"
MODEL="${MODEL:-llama-3.1-405b}"
xargs -a "$HOSTS" -I{} ssh {} tmux new-session -d -s "$SESSION" -c "$(pwd)" \
  -e HF_HOME="${HF_HOME:-/home/ubuntu/.cache/huggingface}" \
  -e OMP_NUM_THREADS=26 \
  -e NCCL_CROSS_NIC=1 \
  -e TORCH_NCCL_AVOID_RECORD_STREAMS=1 \
  -e WANDB_MODE="${WANDB_MODE:-offline}" \
  python -m torch.distributed.run \
    --rdzv-id "$EXPERIMENT" --rdzv-backend c10d --rdzv-endpoint "$HEAD:5001" \
    --nnodes "$NNODES" --nproc-per-node 8 \
    --redirects 3 --log-dir ../logs \
    train_llm.py -e "$EXPERIMENT" -d synthetic -m "$MODEL" \
    -b 1 -s 4096 --cpu-offload on --log-freq 1
