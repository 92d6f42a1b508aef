#!/bin/bash
# Whole GPU suite as the driver runs it, then the no-alloc test alone in a fresh process, then smoke.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider > gpurun_out/gpu_suite.log 2>&1
echo "=== full -m gpu suite exit $?"; tail -n 6 gpurun_out/gpu_suite.log | cut -c1-300
timeout 600 python -m pytest tests/test_model_gpu.py -q -m gpu -s -p no:cacheprovider 2>&1 | grep -E "oracle|stem:" | cut -c1-250
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -p no:cacheprovider > gpurun_out/zz_alone.log 2>&1
echo "=== zz alone exit $?"; tail -n 3 gpurun_out/zz_alone.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo "=== smoke exit $?"; tail -n 2 gpurun_out/smoke.log
