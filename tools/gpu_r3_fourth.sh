#!/bin/bash
# Round-3 fourth GPU pass: (msub, nprod) matrix for the pair im2col layers, stem forms, the two-lane JPEG path.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
B2I_STEM=2 timeout 600 python -m pytest tests/test_model_gpu.py -q -m gpu -p no:cacheprovider -s -k "fused_stem_matches or resnet18_fixture or variants or jpeg or bad_image or member_predict or predict_raw" > gpurun_out/suite_stem2.log 2>&1
echo "=== stem / raw / jpeg tests, B2I_STEM=2: exit $?"; grep -E "passed|failed|error|stem:" gpurun_out/suite_stem2.log | tail -n 6 | cut -c1-300; grep -E "^E  |^FAILED" gpurun_out/suite_stem2.log | head -10 | cut -c1-300
run() {  # label, env...
  label=$1; shift
  env "$@" timeout 300 python bench.py --steps 20 --warmup 5 --no-extras --cpu-budget 0.3 > gpurun_out/ab_$label.json 2> gpurun_out/ab_$label.err || tail -2 gpurun_out/ab_$label.err
  python - "$label" <<P
import json,sys
try:
    d=json.load(open('gpurun_out/ab_%s.json' % sys.argv[1]))
    lm=d['launch_ms']
    pick=['conv:stem(normalize+conv1+maxpool)','conv:layer2.0.conv1+downsample','conv:layer3.0.conv1+downsample','conv:layer3.0.conv2','conv:layer3.1.conv1','conv:layer4.0.conv1+downsample','conv:layer4.0.conv2','conv:layer4.1.conv1','fc:fc']
    print('%-18s value %6d e2e %6d ms/step %.4f | ' % (sys.argv[1], d['value'], d['e2e']['value'], d['ms_per_step']) + ' '.join('%.4f' % lm[k] for k in pick))
except Exception as e: print(sys.argv[1], 'no json', e)
P
}
echo "columns: stem l2.0c1+ds l3.0c1+ds l3.0c2 l3.1c1 l4.0c1+ds l4.0c2 l4.1c1 fc"
for ms in auto 1 2; do for np in 1 2 3 4; do
  if [ $ms = auto ]; then run ms${ms}_np$np B2I_NPROD=$np; else run ms${ms}_np$np B2I_NPROD=$np B2I_MSUB=$ms; fi
done; done
run stem2_np2 B2I_STEM=2 B2I_NPROD=2
run stem1_np2 B2I_STEM=1 B2I_NPROD=2
timeout 600 python tools/jpeg_bench.py --reps 3 > gpurun_out/jpeg_bench2.log 2>&1
echo "=== jpeg_bench exit $?"; grep -E "arrangement" gpurun_out/jpeg_bench2.log | cut -c1-200
timeout 300 python - <<'P' 2>&1 | tail -5
# two callers at once on one model: decode of one batch overlaps the GPU part of the other
import os, sys, time, threading
sys.path.insert(0, '.')
import dmlc_b200 as b2
from oracle import synth
root='gpurun_in/imagenet_1k/train'
wn=sorted(os.listdir(root))
blobs=[open(os.path.join(root,w,sorted(os.listdir(os.path.join(root,w)))[0]),'rb').read() for w in wn]
ot=synth.ensure_ot('resnet18','tests/_cache/resnet18.ot')
ctx=b2.Context((0,))
with b2.Model(ctx,'resnet18',ot,max_batch=256) as m:
    chunks=[blobs[i:i+250] for i in range(0,1000,250)]
    for c in chunks: m.predict_jpeg(c,k=5)
    for nthreads in (1,2,3):
        best=0
        for rep in range(3):
            t0=time.perf_counter()
            def work(j):
                for c in chunks[j::nthreads]: m.predict_jpeg(c,k=5)
            th=[threading.Thread(target=work,args=(j,)) for j in range(nthreads)]
            [t.start() for t in th]; [t.join() for t in th]
            best=max(best,1000/(time.perf_counter()-t0))
        print('predict_jpeg, %d caller thread(s): %.0f images/s' % (nthreads,best))
P
