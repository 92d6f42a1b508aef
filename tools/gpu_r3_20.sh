#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
B2I_PLAN_DEBUG=1 timeout 600 python -m pytest tests/test_model_gpu.py tests/test_fullsize_gpu.py tests/test_kernels_gpu.py -q -m gpu -p no:cacheprovider -x > gpurun_out/suite_bn192.log 2>&1
echo "=== model + fullsize + kernel tests: exit $?"; grep -E "passed|failed|error" gpurun_out/suite_bn192.log | tail -n 2; grep -E "^E  |^FAILED" gpurun_out/suite_bn192.log | head -10 | cut -c1-300
grep -h "\[plan\] layer4" gpurun_out/suite_bn192.log | sort | uniq -c | head
run() {  # label, env...
  label=$1; shift
  env "$@" timeout 300 python bench.py --steps 20 --warmup 5 --no-extras --cpu-budget 0.3 > gpurun_out/ab_$label.json 2> gpurun_out/ab_$label.err || tail -2 gpurun_out/ab_$label.err
  python - "$label" <<P
import json,sys
try:
    d=json.load(open('gpurun_out/ab_%s.json' % sys.argv[1]))
    L=d['launch_ms']; g=lambda k: L.get('conv:'+k, 0)
    print('%-14s value %7d ms/step %.4f | l2.0c1 %.4f | l3.0c2 %.4f l3.1c1 %.4f | l4.0c1 %.4f l4.0c2 %.4f l4.1c1 %.4f l4.1c2 %.4f' % (sys.argv[1], d['value'], d['ms_per_step'], g('layer2.0.conv1+downsample'), g('layer3.0.conv2'), g('layer3.1.conv1'), g('layer4.0.conv1+downsample'), g('layer4.0.conv2'), g('layer4.1.conv1'), g('layer4.1.conv2')))
except Exception as e: print(sys.argv[1], 'no json', e)
P
}
run new
run no192 B2I_NO_BN192=1
run new_b
run no192_b B2I_NO_BN192=1

