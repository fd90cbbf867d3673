#!/usr/bin/env bash
# transport comparison at N GPUs (one gpurun call)
cd "$(dirname "$0")/.."; mkdir -p gpurun_out; export PYTHONUNBUFFERED=1
N=${N:-4}; TAG=${TAG:-n4}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
run() { name=$1; shift; echo "=== $name"; timeout -s KILL ${TMO:-400} "$@" > gpurun_out/$name.log 2>&1; echo "rc=$?"; grep -E '^\{' gpurun_out/$name.log | tail -1 | python -c "
import sys, json
for l in sys.stdin:
    try:
        d = json.loads(l); print({k: d[k] for k in ('value', 'ms_per_step', 'n_gpus') if k in d}, d.get('e2e', {}).get('ms_per_step'), d.get('final_loss'))
    except Exception: print(l[:200])
"; }
for s in ${STAGES:-par bench}; do
  case $s in
    par) TMO=900 run ${TAG}_par python -m pytest tests/test_parallel_gpu.py tests/test_peer_gpu.py tests/test_fused_ag_gpu.py -q -k "${PARK:-4 or hsdp or 2d or peer or nvls or transports or multi_gpu}"; tail -8 gpurun_out/${TAG}_par.log ;;
    bench)
      B200TRAIN_RS_TRANSPORT=nvls run ${TAG}_kernel_nvls $TR bench.py --gpus $N --steps 5 --warmup 3 --fuse-optimizer kernel
      B200TRAIN_RS_TRANSPORT=ce run ${TAG}_kernel_ce $TR bench.py --gpus $N --steps 5 --warmup 3 --fuse-optimizer kernel
      B200TRAIN_RS_TRANSPORT=nvls run ${TAG}_stream_nvls $TR bench.py --gpus $N --steps 5 --warmup 3 --fuse-optimizer stream ;;
    ref) run ${TAG}_ref $TR bench.py --impl reference --gpus $N --steps 5 --warmup 3 ;;
    tl) run ${TAG}_tl $TR bench/step_timeline.py --out gpurun_out/timeline_w${N}_${TAG}.json --windows 3 --fuse kernel
        grep -E "^---|kernel|Memcpy|us at" gpurun_out/${TAG}_tl.log | head -50 ;;
  esac
done
