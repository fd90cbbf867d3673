#!/usr/bin/env bash
# 2-GPU check of the sharded checkpoint exchange over NCCL + peer memory: 4 steps with a checkpoint at step 4,
# resume for 2 more, against 6 uninterrupted steps (losses of steps 5 and 6 must agree); then the peer-memory
# collective tests (the two-phase all-reduce now ends with an exit barrier).
set -u
cd "$(dirname "$0")/../04-fully-sharded-data-parallel"
OUT=/tmp/r2_ckpt2; rm -rf $OUT
RUN="python -m torch.distributed.run --standalone --local-addr 127.0.0.1 --nproc-per-node 2 train_llm.py -d synthetic:64 -m llama-smoke -b 2 -s 256 --ckpt-freq 4 --log-freq 1 --save-dir $OUT --progress-bar off --wandb-mode off --lr 1e-3"
$RUN -e a --max-steps 4 2>&1 | grep -o "rank=0.*running_loss': [0-9.]*" | grep -o "global_step': [0-9]*\|running_loss': [0-9.]*" | paste - - 
ls $OUT/a $OUT/a/checkpoint | tr '\n' ' '; echo
$RUN -e a --max-steps 2 2>&1 | grep -o "rank=0.*running_loss': [0-9.]*" | grep -o "global_step': [0-9]*\|running_loss': [0-9.]*" | paste - -
echo "--- uninterrupted"
$RUN -e b --max-steps 6 2>&1 | grep -o "rank=0.*running_loss': [0-9.]*" | grep -o "global_step': [0-9]*\|running_loss': [0-9.]*" | paste - -
cd ..
timeout 150 python -m pytest tests/test_peer_gpu.py -x -q -m gpu 2>&1 | tail -3
