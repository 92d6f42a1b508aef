#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() {  # label, env...
  label=$1; shift
  env "$@" timeout 300 python bench.py --steps 20 --warmup 5 --no-extras --cpu-budget 0.3 > gpurun_out/ab_$label.json 2> gpurun_out/ab_$label.err || tail -2 gpurun_out/ab_$label.err
  python - "$label" <<P
import json,sys
try:
    d=json.load(open('gpurun_out/ab_%s.json' % sys.argv[1]))
    L=d['launch_ms']; g=lambda k: L.get('conv:'+k, 0)
    print('%-14s value %7d ms/step %.4f | l1 %.4f %.4f %.4f %.4f | l4 %.4f %.4f %.4f' % (sys.argv[1], d['value'], d['ms_per_step'], g('layer1.0.conv1'), g('layer1.0.conv2'), g('layer1.1.conv1'), g('layer1.1.conv2'), g('layer4.0.conv2'), g('layer4.1.conv1'), g('layer4.1.conv2')))
except Exception as e: print(sys.argv[1], 'no json', e)
P
}
run np1 B2I_PATCH_NPROD=1
run np2 B2I_PATCH_NPROD=2
run np3 B2I_PATCH_NPROD=3
run np1_b B2I_PATCH_NPROD=1
run np3_b B2I_PATCH_NPROD=3
B2I_PATCH_NPROD=3 timeout 600 python -m pytest tests/test_model_gpu.py tests/test_kernels_gpu.py -q -m gpu -p no:cacheprovider -x 2>&1 | tail -2
