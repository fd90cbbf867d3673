#!/usr/bin/env bash
# One gpurun call: kernel tests, kernel microbench, end-to-end bench. Each stage runs under its own
# timeout and logs to gpurun_out/; a failing stage does not stop the later ones.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
stage() {  # name timeout cmd...
  local name=$1 t=$2; shift 2
  echo "=== $name ($(date +%T))"
  timeout -s KILL "$t" "$@" > "gpurun_out/$name.log" 2>&1
  local rc=$?
  echo "$name rc=$rc"
  tail -n "${TAIL:-6}" "gpurun_out/$name.log"
}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.used --format=csv > gpurun_out/00_smi.log 2>&1
STAGES="${STAGES:-sanity gemm elem attn model kbench bench ref}"
for s in $STAGES; do
  case $s in
    sanity) stage 01_gemm_sanity 120 python -m pytest tests/test_gemm_gpu.py -q -x -k "shape0 or shape1" ;;
    gemm)   stage 02_gemm 600 python -m pytest tests/test_gemm_gpu.py -q ;;
    elem)   stage 03_elementwise 600 python -m pytest tests/test_elementwise_gpu.py -q ;;
    attn)   stage 04_attention 600 python -m pytest tests/test_attention_gpu.py -q ;;
    model)  stage 05_model 600 python -m pytest tests/test_model_gpu.py -q ;;
    kbench) TAIL=40 stage 06_kernel_bench 900 python bench/kernel_bench.py ;;
    bench)  stage 07_bench 900 python bench.py --steps ${BENCH_STEPS:-5} --warmup 3 ;;
    ref)    stage 08_bench_ref 1200 python bench.py --impl reference --steps ${BENCH_STEPS:-5} --warmup 3 ;;
    all)    stage 09_pytest_all 1500 python -m pytest tests -q -m gpu ;;
  esac
done
echo "=== done ($(date +%T))"
