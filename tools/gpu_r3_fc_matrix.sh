#!/bin/bash
# FC plan matrix on AlexNet: tile width x split-K factor x A-box rows, batch 32 and 256 (per-launch ms of the three FC layers).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_model_gpu.py -q -m gpu -p no:cacheprovider -x -k "variants or ragged" 2>&1 | tail -2
run() {  # label, batch, env...
  label=$1; batch=$2; shift 2
  env "$@" timeout 200 python bench.py --model alexnet --batch $batch --steps 20 --warmup 5 --no-extras --cpu-budget 0.2 > gpurun_out/fc_$label.json 2> gpurun_out/fc_$label.err || tail -2 gpurun_out/fc_$label.err
  python - "$label" <<P
import json,sys
try:
    d=json.load(open('gpurun_out/fc_%s.json' % sys.argv[1])); lm=d['launch_ms']
    print('%-26s ms/step %.4f | c1 %.4f c4 %.4f c6 %.4f' % (sys.argv[1], d['ms_per_step'], lm['fc:classifier.1'], lm['fc:classifier.4'], lm['fc:classifier.6']))
except Exception as e: print(sys.argv[1], 'no json', e)
P
}
for batch in 32 256; do
  run b${batch}_old $batch B2I_NO_SPLITK=1
  for bn in 64 128 256; do for sp in 1 2 4 8; do
    run b${batch}_bn${bn}_sp${sp}_a128 $batch B2I_FC_BN=$bn B2I_FC_SPLITS=$sp B2I_FC_AROWS=128
    if [ $batch = 32 ]; then run b${batch}_bn${bn}_sp${sp}_a32 $batch B2I_FC_BN=$bn B2I_FC_SPLITS=$sp B2I_FC_AROWS=32; fi
  done; done
done
