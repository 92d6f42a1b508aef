#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/ -q -m gpu -p no:cacheprovider > gpurun_out/gpu_suite.log 2>&1
echo "=== full -m gpu suite exit $?"; grep -E "passed|failed|error" gpurun_out/gpu_suite.log | tail -n 3 | cut -c1-300
grep -E "^E  |^FAILED" gpurun_out/gpu_suite.log | head -20 | cut -c1-400
for i in 1 2 3 4 5 6 7 8; do timeout 300 python -m pytest tests/test_model_gpu.py -q -m gpu -p no:cacheprovider -k "jpeg" 2>&1 | grep -E "passed|failed|differ" | cut -c1-300; done
run() {  # label, env...
  label=$1; shift
  env "$@" timeout 300 python bench.py --steps 20 --warmup 5 --no-extras --cpu-budget 0.3 > gpurun_out/ab_$label.json 2> gpurun_out/ab_$label.err || tail -2 gpurun_out/ab_$label.err
  python - "$label" <<P
import json,sys
try:
    d=json.load(open('gpurun_out/ab_%s.json' % sys.argv[1]))
    print('%-12s value %7d e2e %7d ms/step %.4f | ' % (sys.argv[1], d['value'], d['e2e']['value'], d['ms_per_step']) + ' '.join('%s=%.4f' % (k.split(':')[-1][:12], v) for k, v in d['launch_ms'].items()))
except Exception as e: print(sys.argv[1], 'no json', e)
P
}
run stem2_a B2I_STEM=2
run stem1_a B2I_STEM=1
run stem2_b B2I_STEM=2
run stem1_b B2I_STEM=1
