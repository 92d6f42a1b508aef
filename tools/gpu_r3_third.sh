#!/bin/bash
# Round-3 third GPU pass: multi-issuer TMA producers (B2I_NPROD) and the second stem form (B2I_STEM=2): parity, then A/B timing.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py -q -m gpu -p no:cacheprovider -x > gpurun_out/suite_nprod4.log 2>&1
echo "=== kernels + model tests, NPROD default(4), stem 1: exit $?"; grep -E "passed|failed|error" gpurun_out/suite_nprod4.log | tail -n 2 | cut -c1-300; grep -E "^E  |^FAILED" gpurun_out/suite_nprod4.log | head -10 | cut -c1-300
B2I_STEM=2 timeout 600 python -m pytest tests/test_model_gpu.py -q -m gpu -p no:cacheprovider -s -k "fused_stem_matches or resnet18_fixture or variants" > gpurun_out/suite_stem2.log 2>&1
echo "=== stem tests, B2I_STEM=2: exit $?"; grep -E "passed|failed|error|stem:" gpurun_out/suite_stem2.log | tail -n 6 | cut -c1-300; grep -E "^E  |^FAILED" gpurun_out/suite_stem2.log | head -10 | cut -c1-300
for cfg in "1 1" "1 2" "1 4" "2 4"; do
  set -- $cfg
  B2I_STEM=$1 B2I_NPROD=$2 timeout 300 python bench.py --steps 20 --warmup 5 --no-extras --cpu-budget 0.5 > gpurun_out/ab_stem$1_nprod$2.json 2> gpurun_out/ab_stem$1_nprod$2.err
  echo "=== bench stem=$1 nprod=$2 exit $?"; tail -2 gpurun_out/ab_stem$1_nprod$2.err | cut -c1-300
  python - <<P
import json
try:
    d=json.load(open('gpurun_out/ab_stem$1_nprod$2.json'))
    print('value',round(d['value']),'e2e',round(d['e2e']['value']),'ms/step',round(d['ms_per_step'],4))
    print({k.replace('conv:',''):v for k,v in d['launch_ms'].items()})
except Exception as e: print('no json',e)
P
done
