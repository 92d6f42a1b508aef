#!/bin/bash
# Round-3 seventh GPU pass: whole suite with the new defaults (stem2, FC plan), FC plan A/B, stem2 ncu.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/ -q -m gpu -p no:cacheprovider > gpurun_out/gpu_suite.log 2>&1
echo "=== full -m gpu suite exit $?"; grep -E "passed|failed|error" gpurun_out/gpu_suite.log | tail -n 3 | cut -c1-300
grep -E "^E  |^FAILED" gpurun_out/gpu_suite.log | head -20 | cut -c1-300
run() {  # label, model, batch, env...
  label=$1; model=$2; batch=$3; shift 3
  env "$@" timeout 300 python bench.py --model $model --batch $batch --steps 20 --warmup 5 --no-extras --cpu-budget 0.3 > gpurun_out/ab_$label.json 2> gpurun_out/ab_$label.err || tail -2 gpurun_out/ab_$label.err
  grep "\[plan\] classifier\|\[plan\] fc" gpurun_out/ab_$label.err | sort -u | cut -c1-160
  python - "$label" <<P
import json,sys
try:
    d=json.load(open('gpurun_out/ab_%s.json' % sys.argv[1]))
    print('%-22s value %7d e2e %7d ms/step %.4f | ' % (sys.argv[1], d['value'], d['e2e']['value'], d['ms_per_step']) + ' '.join('%s=%.4f' % (k.split(':')[-1][:14], v) for k, v in d['launch_ms'].items()))
except Exception as e: print(sys.argv[1], 'no json', e)
P
}
run alex32_plan alexnet 32 B2I_PLAN_DEBUG=1
run alex32_noplan alexnet 32 B2I_NO_SPLITK=1
run alex256_plan alexnet 256 B2I_PLAN_DEBUG=1
run alex256_noplan alexnet 256 B2I_NO_SPLITK=1
run r18_plan resnet18 256 B2I_PLAN_DEBUG=1
run r18_noplan resnet18 256 B2I_NO_SPLITK=1
python tools/profile_step.py 2 1 > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:stem2_kernel -s 2 -c 1 -f -o gpurun_out/prof_stem2b python tools/profile_step.py 2 1 > gpurun_out/ncu_stem2b.log 2>&1
echo "stem2 ncu exit $?"; tail -2 gpurun_out/ncu_stem2b.log
