#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
B2I_PAIR_SPLITS="256:2,512:3" timeout 600 python -m pytest tests/test_model_gpu.py tests/test_fullsize_gpu.py -q -m gpu -p no:cacheprovider -x > gpurun_out/suite_split.log 2>&1
echo "=== model + fullsize tests, pair splits 256:2,512:3: exit $?"; grep -E "passed|failed|error" gpurun_out/suite_split.log | tail -n 2; grep -E "^E  |^FAILED" gpurun_out/suite_split.log | head -10 | cut -c1-300
run() {  # label, env...
  label=$1; shift
  env "$@" timeout 300 python bench.py --steps 20 --warmup 5 --no-extras --cpu-budget 0.3 > gpurun_out/ab_$label.json 2> gpurun_out/ab_$label.err || tail -2 gpurun_out/ab_$label.err
  python - "$label" <<P
import json,sys
try:
    d=json.load(open('gpurun_out/ab_%s.json' % sys.argv[1]))
    L=d['launch_ms']; g=lambda k: L.get('conv:'+k, 0)
    print('%-14s value %7d ms/step %.4f | l3.0c2 %.4f l3.1c1 %.4f l3.1c2 %.4f | l4.0c1 %.4f l4.0c2 %.4f l4.1c1 %.4f l4.1c2 %.4f' % (sys.argv[1], d['value'], d['ms_per_step'], g('layer3.0.conv2'), g('layer3.1.conv1'), g('layer3.1.conv2'), g('layer4.0.conv1+downsample'), g('layer4.0.conv2'), g('layer4.1.conv1'), g('layer4.1.conv2')))
except Exception as e: print(sys.argv[1], 'no json', e)
P
}
run base
run l4s2m1 B2I_PAIR_SPLITS=512:2 B2I_MSUB=1
run l4s3m1 B2I_PAIR_SPLITS=512:3 B2I_MSUB=1
run l4s4m1 B2I_PAIR_SPLITS=512:4 B2I_MSUB=1
run l4s3 B2I_PAIR_SPLITS=512:3
run l4s2 B2I_PAIR_SPLITS=512:2
run l3s2 B2I_PAIR_SPLITS=256:2
run l3s3 B2I_PAIR_SPLITS=256:3
run l3s4 B2I_PAIR_SPLITS=256:4
run base_m1 B2I_MSUB=1
run base2
