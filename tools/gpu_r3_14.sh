#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() {  # label, args...
  label=$1; shift
  B2I_PLAN_DEBUG=1 timeout 300 python bench.py --steps 20 --warmup 5 --no-extras --cpu-budget 0.3 "$@" > gpurun_out/ab_$label.json 2> gpurun_out/ab_$label.err || tail -2 gpurun_out/ab_$label.err
  grep "\[plan\] layer1.0.conv1" gpurun_out/ab_$label.err | sort -u | head -2
  python - "$label" <<P
import json,sys
try:
    d=json.load(open('gpurun_out/ab_%s.json' % sys.argv[1]))
    print('%-12s value %7d e2e %7d ms/step %.4f | ' % (sys.argv[1], d['value'], d['e2e']['value'], d['ms_per_step']) + ' '.join('%s=%.4f' % (k.split(':')[-1][:12], v) for k, v in list(d['launch_ms'].items())[:7]))
except Exception as e: print(sys.argv[1], 'no json', e)
P
}
run patch
run flat --patch-conv 0
run patch2
run flat2 --patch-conv 0
