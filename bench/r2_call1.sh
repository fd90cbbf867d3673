#!/usr/bin/env bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out; export PYTHONUNBUFFERED=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
run() { name=$1; shift; echo "=== $name"; timeout -s KILL 300 "$@" > gpurun_out/$name.log 2>&1; echo "rc=$?"; grep -E '^\{' gpurun_out/$name.log | tail -1 | cut -c1-400; }
run r2c1_tl1 python bench/step_timeline.py --out gpurun_out/timeline_w1.json
run r2c1_tl2 $TR bench/step_timeline.py --out gpurun_out/timeline_w2.json
run r2c1_b2_default $TR bench.py --gpus 2 --steps 6 --warmup 3
B200TRAIN_RS_BLOCKS=128 run r2c1_b2_rs128 $TR bench.py --gpus 2 --steps 6 --warmup 3
B200TRAIN_RS_BLOCKS=32 run r2c1_b2_rs32 $TR bench.py --gpus 2 --steps 6 --warmup 3
B200TRAIN_NVLS=1 run r2c1_b2_nvls $TR bench.py --gpus 2 --steps 6 --warmup 3
run r2c1_b2_kernel $TR bench.py --gpus 2 --steps 6 --warmup 3 --fuse-optimizer kernel
run r2c1_b2_off $TR bench.py --gpus 2 --steps 6 --warmup 3 --fuse-optimizer off
B200TRAIN_AG=kernel run r2c1_b2_agk $TR bench.py --gpus 2 --steps 6 --warmup 3
run r2c1_b1 python bench.py --steps 6 --warmup 3
