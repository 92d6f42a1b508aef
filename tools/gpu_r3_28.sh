#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for b in 32 64 128 256 512; do
  timeout 300 python bench.py --batch $b --steps 20 --warmup 5 --no-extras --cpu-budget 0.2 > gpurun_out/bs_$b.json 2> gpurun_out/bs_$b.err || tail -2 gpurun_out/bs_$b.err
  python - $b <<'P'
import json,sys
b=int(sys.argv[1])
d=json.load(open('gpurun_out/bs_%d.json' % b))
L=d['launch_ms']
print('bs %4d step %.4f ms | per 256 images:' % (b, d['ms_per_step']), ' '.join('%s=%.4f' % (k.replace('conv:','').replace('layer','l').replace('.conv','c').replace('+downsample','+d')[:10], v*256.0/b) for k,v in L.items()))
P
done
