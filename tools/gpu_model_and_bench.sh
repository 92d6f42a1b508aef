#!/bin/bash
# Model-level parity tests, smoke and a short bench on one B200; logs under gpurun_out/.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_model_gpu.py -m gpu -q -s -p no:cacheprovider > gpurun_out/model_tests.log 2>&1
echo "=== model tests exit $?"
grep -E "gap|max-abs/scale|passed|failed|FAILED|Error" gpurun_out/model_tests.log | cut -c1-300 | tail -60
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo "=== smoke exit $?"; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 --profile-out gpurun_out/launch_share.csv > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "=== bench exit $?"; tail -c 3000 gpurun_out/bench.json; tail -5 gpurun_out/bench.err
