#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_model_gpu.py tests/test_fullsize_gpu.py tests/test_kernels_gpu.py -q -m gpu -p no:cacheprovider -x 2>&1 | tail -3
run() {  # label, env...
  label=$1; shift
  env "$@" timeout 300 python bench.py --steps 20 --warmup 5 --no-extras --cpu-budget 0.2 > gpurun_out/v_$label.json 2> gpurun_out/v_$label.err || tail -2 gpurun_out/v_$label.err
  python - "$label" <<'P'
import json,sys
d=json.load(open('gpurun_out/v_%s.json' % sys.argv[1]))
L=d['launch_ms']
print('%-10s value %d step %.4f launches %d |' % (sys.argv[1], d['value'], d['ms_per_step'], d['gpu_launches']), ' '.join('%s=%.4f' % (k.replace('conv:','').replace('layer','l').replace('.conv','c').replace('+downsample','+d')[:9], v) for k,v in L.items() if 'layer4' in k or 'layer3.1' in k))
P
}
run tail A=1
run notail B2I_NO_TAILSPLIT=1
run tail_b A=1
run notail_b B2I_NO_TAILSPLIT=1
