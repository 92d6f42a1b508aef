#!/bin/bash
# Round-3 sixth GPU pass: whole suite, stem forms A/B, split-K A/B on AlexNet, two-caller JPEG path.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/ -q -m gpu -p no:cacheprovider > gpurun_out/gpu_suite.log 2>&1
echo "=== full -m gpu suite exit $?"; grep -E "passed|failed|error" gpurun_out/gpu_suite.log | tail -n 3 | cut -c1-300
grep -E "^E  |^FAILED" gpurun_out/gpu_suite.log | head -30 | cut -c1-300
B2I_STEM=2 timeout 600 python -m pytest tests/test_model_gpu.py tests/test_fullsize_gpu.py -q -m gpu -p no:cacheprovider > gpurun_out/suite_stem2.log 2>&1
echo "=== model + fullsize tests, B2I_STEM=2: exit $?"; grep -E "passed|failed|error" gpurun_out/suite_stem2.log | tail -n 2; grep -E "^E  |^FAILED" gpurun_out/suite_stem2.log | head -10 | cut -c1-300
run() {  # label, model, batch, env...
  label=$1; model=$2; batch=$3; shift 3
  env "$@" timeout 300 python bench.py --model $model --batch $batch --steps 20 --warmup 5 --no-extras --cpu-budget 0.3 > gpurun_out/ab_$label.json 2> gpurun_out/ab_$label.err || tail -2 gpurun_out/ab_$label.err
  python - "$label" <<P
import json,sys
try:
    d=json.load(open('gpurun_out/ab_%s.json' % sys.argv[1]))
    print('%-22s value %7d e2e %7d ms/step %.4f | ' % (sys.argv[1], d['value'], d['e2e']['value'], d['ms_per_step']) + ' '.join('%s=%.4f' % (k.split(':')[-1][:14], v) for k, v in d['launch_ms'].items()))
except Exception as e: print(sys.argv[1], 'no json', e)
P
}
run r18_stem1 resnet18 256 B2I_STEM=1
run r18_stem2 resnet18 256 B2I_STEM=2
run r18_stem1_b resnet18 256 B2I_STEM=1
run r18_stem2_b resnet18 256 B2I_STEM=2
run alex32_splitk alexnet 32 B2I_X=0
run alex32_nosplit alexnet 32 B2I_NO_SPLITK=1
run alex256_splitk alexnet 256 B2I_X=0
run alex256_nosplit alexnet 256 B2I_NO_SPLITK=1
run r18_nosplit resnet18 256 B2I_NO_SPLITK=1
timeout 300 python - <<'P' 2>&1 | tail -6
import os, sys, time, threading
sys.path.insert(0, '.')
import dmlc_b200 as b2
from oracle import synth
root='gpurun_in/imagenet_1k/train'
wn=sorted(os.listdir(root))
blobs=[open(os.path.join(root,w,sorted(os.listdir(os.path.join(root,w)))[0]),'rb').read() for w in wn]
ot=synth.ensure_ot('resnet18','tests/_cache/resnet18.ot')
ctx=b2.Context((0,))
print('backend', ctx.jpeg_backend())
with b2.Model(ctx,'resnet18',ot,max_batch=256) as m:
    chunks=[blobs[i:i+250] for i in range(0,1000,250)]
    for c in chunks: m.predict_jpeg(c,k=5)
    for c in chunks: m.predict_jpeg(c,k=5)
    for nthreads in (1,2,3):
        best=0
        for rep in range(4):
            t0=time.perf_counter()
            def work(j):
                for c in chunks[j::nthreads]: m.predict_jpeg(c,k=5)
            th=[threading.Thread(target=work,args=(j,)) for j in range(nthreads)]
            [t.start() for t in th]; [t.join() for t in th]
            best=max(best,1000/(time.perf_counter()-t0))
        print('predict_jpeg, %d caller thread(s): %.0f images/s' % (nthreads,best))
P
