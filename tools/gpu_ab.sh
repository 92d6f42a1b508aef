#!/bin/bash
# A/B inside one GPU session, alternating runs.  Default: ab/lib_old.so (B2I_LIB) vs the in-tree library.
# AB_OLD_ENV="B2I_NO_PDL=1" compares an environment switch of the in-tree library instead.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
OLD_ENV=${AB_OLD_ENV:-B2I_LIB=$PWD/ab/lib_old.so}
for r in 1 2; do
  for v in old new; do
    if [ $v = old ]; then E="$OLD_ENV"; else E="B2I_AB=new"; fi
    env $E timeout 300 python bench.py --steps 30 --warmup 5 --cpu-budget 0.5 $BENCH_ARGS > gpurun_out/ab_${v}_$r.json 2> gpurun_out/ab_${v}_$r.err
    python -c "
import json,sys; d=json.load(open('gpurun_out/ab_${v}_$r.json'))
print('$v $r value',round(d['value']),'e2e',round(d['e2e']['value']),'ms',round(d['ms_per_step'],4),'stem',d['launch_ms'].get('conv:stem(normalize+conv1+maxpool)'))" || tail -3 gpurun_out/ab_${v}_$r.err
  done
done
