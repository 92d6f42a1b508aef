#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() {  # label, env...
  label=$1; shift
  env "$@" timeout 300 python bench.py --steps 20 --warmup 5 --no-extras --cpu-budget 0.2 > gpurun_out/v_$label.json 2> gpurun_out/v_$label.err || tail -2 gpurun_out/v_$label.err
  python - "$label" <<'P'
import json,sys
d=json.load(open('gpurun_out/v_%s.json' % sys.argv[1]))
L=d['launch_ms']
print('%-10s value %d step %.4f |' % (sys.argv[1], d['value'], d['ms_per_step']), ' '.join('%s=%.4f' % (k.replace('conv:','').replace('layer','l').replace('.conv','c').replace('+downsample','+d')[:9], v) for k,v in L.items() if 'layer1' in k or 'layer2.0' in k))
P
}
run acc4 A=1
run old B2I_LIB=/root/repo/gpurun_in/libb200infer_prev.so
run acc4_b A=1
timeout 600 python -m pytest tests/test_model_gpu.py tests/test_kernels_gpu.py -q -m gpu -p no:cacheprovider -x 2>&1 | tail -2
