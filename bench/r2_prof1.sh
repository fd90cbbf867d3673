#!/usr/bin/env bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out; export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_parity_reference_gpu.py -q -s > gpurun_out/parity.log 2>&1; echo "parity rc=$?"; grep -E "ours|reference|passed|failed|Error" gpurun_out/parity.log | tail -6
python bench/prof_kernels.py gemm_fused > gpurun_out/prof_plain_fused.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gemm_bf16_kernel -s 4 -c 5 -o gpurun_out/prof_gemm_fused python bench/prof_kernels.py gemm_fused > gpurun_out/ncu_gemm_fused.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_gemm_fused.log
