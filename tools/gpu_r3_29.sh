#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() {  # label, extra args, env...
  label=$1; shift; extra=$1; shift
  env "$@" timeout 300 python bench.py --steps 20 --warmup 5 --no-extras --cpu-budget 0.2 $extra > gpurun_out/v_$label.json 2> gpurun_out/v_$label.err || tail -2 gpurun_out/v_$label.err
  python - "$label" <<'P'
import json,sys
d=json.load(open('gpurun_out/v_%s.json' % sys.argv[1]))
L=d['launch_ms']
print('%-14s step %.4f |' % (sys.argv[1], d['ms_per_step']), ' '.join('%s=%.4f' % (k.replace('conv:','').replace('layer','l').replace('.conv','c').replace('+downsample','+d').replace('.downsample','ds')[:9], v) for k,v in L.items() if ('.0.' in k and ('conv1' in k or 'downsample' in k))))
P
}
run base "" A=1
run nofuse "--fuse-downsample 0" A=1
run noresb "" B2I_NO_RESB=1
run msub4 "" B2I_MSUB=4
run msub1 "" B2I_MSUB=1
run nosnake "" B2I_NO_SNAKE=1
run nopdl "" B2I_NO_PDL=1
