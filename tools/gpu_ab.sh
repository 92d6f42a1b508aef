#!/bin/bash
# A/B of two builds inside one GPU session: ab/lib_old.so vs the in-tree library, alternating runs.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for r in 1 2; do
  for v in old new; do
    if [ $v = old ]; then export B2I_LIB=$PWD/ab/lib_old.so; else unset B2I_LIB; fi
    timeout 300 python bench.py --steps 30 --warmup 5 --cpu-budget 0.5 $BENCH_ARGS > gpurun_out/ab_${v}_$r.json 2> gpurun_out/ab_${v}_$r.err
    python -c "
import json,sys; d=json.load(open('gpurun_out/ab_${v}_$r.json'))
print('$v $r value',round(d['value']),'e2e',round(d['e2e']['value']),'ms',round(d['ms_per_step'],4),'stem',d['launch_ms'].get('conv:stem(normalize+conv1+maxpool)'))" || tail -3 gpurun_out/ab_${v}_$r.err
  done
done
