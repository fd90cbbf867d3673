#!/usr/bin/env bash
# One compute-sanitizer tool per invocation (memcheck | racecheck | synccheck | initcheck) over the
# small-shape kernel driver; the summary goes to gpurun_out/sanitizer_<tool>.log.
#   bash bench/sanitize.sh memcheck [gemm|attn|elementwise|all]
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TOOL=${1:-memcheck}
WHICH=${2:-all}
timeout -s KILL ${SAN_TIMEOUT:-400} python bench/sanitize_driver.py "$WHICH" > gpurun_out/sanitize_plain.log 2>&1 || { echo "driver failed without the tool"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
timeout -s KILL ${SAN_TIMEOUT:-400} compute-sanitizer --tool "$TOOL" --print-limit 20 \
    python bench/sanitize_driver.py "$WHICH" > "gpurun_out/sanitizer_${TOOL}_${WHICH}.log" 2>&1
echo "rc=$?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|ok$|done" "gpurun_out/sanitizer_${TOOL}_${WHICH}.log" | tail -8
