#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py tests/test_fullsize_gpu.py -q -m gpu -p no:cacheprovider > gpurun_out/gpu_suite.log 2>&1
echo "=== kernels+model+fullsize exit $?"; grep -E "passed|failed|error" gpurun_out/gpu_suite.log | tail -n 3 | cut -c1-300
grep -E "^E  |^FAILED" gpurun_out/gpu_suite.log | head -20 | cut -c1-300
run() {  # label, model, batch, env...
  label=$1; model=$2; batch=$3; shift 3
  env "$@" timeout 300 python bench.py --model $model --batch $batch --steps 20 --warmup 5 --no-extras --cpu-budget 0.3 > gpurun_out/ab_$label.json 2> gpurun_out/ab_$label.err || tail -2 gpurun_out/ab_$label.err
  python - "$label" <<P
import json,sys
try:
    d=json.load(open('gpurun_out/ab_%s.json' % sys.argv[1]))
    print('%-22s value %7d e2e %7d ms/step %.4f | ' % (sys.argv[1], d['value'], d['e2e']['value'], d['ms_per_step']) + ' '.join('%s=%.4f' % (k.split(':')[-1][:14], v) for k, v in d['launch_ms'].items() if 'classifier' in k or k.startswith('fc') or 'features' in k))
except Exception as e: print(sys.argv[1], 'no json', e)
P
}
for b in 1 32 256; do
for np in 1 2 3; do run alex${b}_np$np alexnet $b B2I_NPROD=$np; done
run alex${b}_np3_nosplit alexnet $b B2I_NPROD=3 B2I_NO_SPLITK=1
done
run r18_np3 resnet18 256 B2I_NPROD=3
run r18_np1 resnet18 256 B2I_NPROD=1
