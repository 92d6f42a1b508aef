#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
B2I_STEM=3 timeout 600 python -m pytest tests/test_model_gpu.py tests/test_fullsize_gpu.py -q -m gpu -p no:cacheprovider -x > gpurun_out/suite_stem3.log 2>&1
echo "=== model + fullsize tests, B2I_STEM=3: exit $?"; grep -E "passed|failed|error" gpurun_out/suite_stem3.log | tail -n 2; grep -E "^E  |^FAILED" gpurun_out/suite_stem3.log | head -10 | cut -c1-300
run() {  # label, env...
  label=$1; shift
  env "$@" timeout 300 python bench.py --steps 20 --warmup 5 --no-extras --cpu-budget 0.3 > gpurun_out/ab_$label.json 2> gpurun_out/ab_$label.err || tail -2 gpurun_out/ab_$label.err
  python - "$label" <<P
import json,sys
try:
    d=json.load(open('gpurun_out/ab_%s.json' % sys.argv[1]))
    print('%-12s value %7d e2e %7d ms/step %.4f | ' % (sys.argv[1], d['value'], d['e2e']['value'], d['ms_per_step']) + ' '.join('%s=%.4f' % (k.split(':')[-1][:12], v) for k, v in list(d['launch_ms'].items())[:3]))
except Exception as e: print(sys.argv[1], 'no json', e)
P
}
run stem2_a B2I_STEM=2
run stem3_a B2I_STEM=3
run stem1_a B2I_STEM=1
run stem2_b B2I_STEM=2
run stem3_b B2I_STEM=3
timeout 300 python - <<'P'
import sys, time
sys.path.insert(0, '.')
import numpy as np
import dmlc_b200 as b2
from oracle import synth
ot=synth.ensure_ot('resnet18','tests/_cache/resnet18.ot')
ctx=b2.Context((0,))
rng=np.random.default_rng(3)
raw=[rng.integers(0,256,(375,500,3),dtype=np.uint8) for _ in range(256)]
with b2.Model(ctx,'resnet18',ot,max_batch=256) as m:
    for _ in range(3): m.predict_raw(raw,k=5)
    ts=[]
    for _ in range(8):
        t0=time.perf_counter(); m.predict_raw(raw,k=5); ts.append(time.perf_counter()-t0)
    print('predict_raw 256 x 375x500: %.2f ms median -> %.0f images/s' % (1000*sorted(ts)[4], 256/sorted(ts)[4]))
P
