#!/bin/bash
# Round-3 second GPU pass: whole suite (all failures shown), decoder arrangements, TMA two-issuer microbenchmark.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/ -q -m gpu -p no:cacheprovider > gpurun_out/gpu_suite.log 2>&1
echo "=== full -m gpu suite exit $?"; grep -E "passed|failed|error" gpurun_out/gpu_suite.log | tail -n 4 | cut -c1-300
grep -E "^E  |^FAILED" gpurun_out/gpu_suite.log | head -30 | cut -c1-300
timeout 900 python tools/jpeg_bench.py > gpurun_out/jpeg_bench.log 2>&1
echo "=== jpeg_bench exit $?"; cut -c1-400 gpurun_out/jpeg_bench.log
timeout 300 tools/microbench/tma_rate > gpurun_out/tma_rate.log 2>&1
echo "=== tma_rate exit $?"; cat gpurun_out/tma_rate.log
