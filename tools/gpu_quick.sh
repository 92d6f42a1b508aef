#!/bin/bash
# im2col/linear kernel groups + model tests + short bench; logs under gpurun_out/.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for k in "im2col" "linear"; do
  timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "$k" -p no:cacheprovider > "gpurun_out/kern_${k}.log" 2>&1
  echo "=== group [$k] exit $?"; grep -E "passed|failed" "gpurun_out/kern_${k}.log" | tail -3
  grep -E "AssertionError: |B2IError" "gpurun_out/kern_${k}.log" | cut -c1-300 | sort | uniq -c | sort -rn | head -8
done
timeout 900 python -m pytest tests/test_model_gpu.py -m gpu -q -s -p no:cacheprovider > gpurun_out/model_tests.log 2>&1
echo "=== model tests exit $?"
grep -E "gap|passed|failed|FAILED|Error" gpurun_out/model_tests.log | cut -c1-200 | tail -30
timeout 600 python bench.py --steps 20 --warmup 5 --cpu-budget 1 --profile-out gpurun_out/launch_share.csv $BENCH_ARGS > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "=== bench exit $?"; python -c "
import json; d=json.load(open('gpurun_out/bench.json'))
print('value',round(d['value']),'e2e',round(d['e2e']['value']),'ms/step',round(d['ms_per_step'],3),'roofline',round(d['roofline']['frac'],3), d['clocks'])
print(d['launch_ms'])"; tail -3 gpurun_out/bench.err
