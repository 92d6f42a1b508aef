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
print('%-10s value %d step %.4f |' % (sys.argv[1], d['value'], d['ms_per_step']), ' '.join('%.4f' % v for k,v in list(L.items())[:17]))
P
}
run base "" A=1
run nopatch "--patch-conv 0" A=1
run msub1 "" B2I_MSUB=1
run msub4 "" B2I_MSUB=4
run stem1 "" B2I_STEM=1
run np2 "" B2I_NPROD=2
run np1 "" B2I_NPROD=1
run base_b "" A=1
