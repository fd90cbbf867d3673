#!/usr/bin/env bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out; export PYTHONUNBUFFERED=1
python bench/prof_comm_local.py > gpurun_out/prof_plain_comm.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"reduce_local_kernel|gemm_bf16_kernel" -c 9 -o gpurun_out/prof_comm_local python bench/prof_comm_local.py > gpurun_out/ncu_comm_local.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_comm_local.log; tail -2 gpurun_out/prof_plain_comm.log
