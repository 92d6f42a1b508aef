#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_model_gpu.py tests/test_fullsize_gpu.py -q -m gpu -p no:cacheprovider -x > gpurun_out/suite_flats.log 2>&1
echo "=== model + fullsize (layer4 streamed flat): exit $?"; grep -E "passed|failed|error" gpurun_out/suite_flats.log | tail -n 2; grep -E "^E  |^FAILED" gpurun_out/suite_flats.log | head -8 | cut -c1-300
B2I_FLATS_MAXW=14 timeout 600 python -m pytest tests/test_model_gpu.py tests/test_fullsize_gpu.py -q -m gpu -p no:cacheprovider -x > gpurun_out/suite_flats14.log 2>&1
echo "=== model + fullsize (layer3 + layer4 streamed flat): exit $?"; grep -E "passed|failed|error" gpurun_out/suite_flats14.log | tail -n 2; grep -E "^E  |^FAILED" gpurun_out/suite_flats14.log | head -8 | cut -c1-300
run() {  # label, env...
  label=$1; shift
  env "$@" timeout 300 python bench.py --steps 20 --warmup 5 --no-extras --cpu-budget 0.2 > gpurun_out/v_$label.json 2> gpurun_out/v_$label.err || tail -2 gpurun_out/v_$label.err
  python - "$label" <<'P'
import json,sys
d=json.load(open('gpurun_out/v_%s.json' % sys.argv[1]))
L=d['launch_ms']
print('%-10s value %d step %.4f |' % (sys.argv[1], d['value'], d['ms_per_step']), ' '.join('%s=%.4f' % (k.replace('conv:','').replace('layer','l').replace('.conv','c').replace('+downsample','+d')[:9], v) for k,v in L.items() if 'layer4' in k or 'layer3' in k))
P
}
run flats7 A=1
run off B2I_FLATS_MAXW=0
run flats14 B2I_FLATS_MAXW=14
run flats7_b A=1
run off_b B2I_FLATS_MAXW=0
