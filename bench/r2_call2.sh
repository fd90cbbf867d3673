#!/usr/bin/env bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out; export PYTHONUNBUFFERED=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
run() { name=$1; shift; echo "=== $name"; timeout -s KILL 300 "$@" > gpurun_out/$name.log 2>&1; echo "rc=$?"; grep -E '^\{' gpurun_out/$name.log | tail -1 | cut -c1-300; }
run r2c2_tl2 $TR bench/step_timeline.py --out gpurun_out/timeline_w2.json --windows 6
run r2c2_tl2cpu $TR bench/step_timeline.py --out gpurun_out/timeline_w2cpu.json --windows 6 --cpu
