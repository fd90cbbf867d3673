#!/usr/bin/env bash
# Multi-GPU scaling investigation (one gpurun call): peer tests, comm microbench, bench variants, timeline.
# usage: N=2 STAGES="peer comm bench tl" bash bench/r2_scale.sh
cd "$(dirname "$0")/.."; mkdir -p gpurun_out; export PYTHONUNBUFFERED=1
N=${N:-2}; TAG=${TAG:-r2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
run() { name=$1; shift; echo "=== $name"; timeout -s KILL ${TMO:-400} "$@" > gpurun_out/$name.log 2>&1; echo "rc=$?"; grep -E '^\{' gpurun_out/$name.log | tail -1 | python -c "
import sys, json
for l in sys.stdin:
    try:
        d = json.loads(l); print({k: d[k] for k in ('value', 'ms_per_step', 'n_gpus') if k in d}, d.get('e2e', {}).get('ms_per_step'), d.get('final_loss'))
    except Exception: print(l[:200])
"; }
for s in ${STAGES:-peer comm bench tl}; do
  case $s in
    peer) run ${TAG}_peer python -m pytest tests/test_peer_gpu.py -x -q; tail -3 gpurun_out/${TAG}_peer.log ;;
    fag)  run ${TAG}_fag python -m pytest tests/test_fused_ag_gpu.py -x -q; tail -3 gpurun_out/${TAG}_fag.log ;;
    par)  TMO=900 run ${TAG}_par python -m pytest tests/test_parallel_gpu.py -q; tail -8 gpurun_out/${TAG}_par.log ;;
    comm) run ${TAG}_comm $TR bench/comm_bench.py; grep -E "rs_adamw|rs_ms|ag_ce|gemm|roofline" gpurun_out/${TAG}_comm.log ;;
    bench)
      run ${TAG}_b_stream $TR bench.py --gpus $N --steps 6 --warmup 3 --fuse-optimizer stream
      run ${TAG}_b_kernel $TR bench.py --gpus $N --steps 6 --warmup 3 --fuse-optimizer kernel
      B200TRAIN_FSDP_FUSED_AG=0 run ${TAG}_b_kernel_nofag $TR bench.py --gpus $N --steps 6 --warmup 3 --fuse-optimizer kernel ;;
    sweep)
      B200TRAIN_RS_TRANSPORT=peer run ${TAG}_b_kernel_peer $TR bench.py --gpus $N --steps 6 --warmup 3 --fuse-optimizer kernel
      B200TRAIN_RS_TRANSPORT=ce run ${TAG}_b_kernel_ce $TR bench.py --gpus $N --steps 6 --warmup 3 --fuse-optimizer kernel
      B200TRAIN_RS_TRANSPORT=ce B200TRAIN_RSADAM_BLOCKS=148 run ${TAG}_b_kernel_ce148 $TR bench.py --gpus $N --steps 6 --warmup 3 --fuse-optimizer kernel
      B200TRAIN_RS_TRANSPORT=nvls run ${TAG}_b_kernel_nvls $TR bench.py --gpus $N --steps 6 --warmup 3 --fuse-optimizer kernel
      B200TRAIN_RS_TRANSPORT=nvls run ${TAG}_b_stream_nvls $TR bench.py --gpus $N --steps 6 --warmup 3 --fuse-optimizer stream
      B200TRAIN_RS_TRANSPORT=ce run ${TAG}_b_stream_ce $TR bench.py --gpus $N --steps 6 --warmup 3 --fuse-optimizer stream ;;
    tl) run ${TAG}_tl $TR bench/step_timeline.py --out gpurun_out/timeline_w${N}_${TAG}.json --windows 4 --fuse kernel
        grep -E "^---|kernel|Memcpy|us at" gpurun_out/${TAG}_tl.log | head -60 ;;
    b1) run ${TAG}_b1 python bench.py --steps 6 --warmup 3 ;;
  esac
done
