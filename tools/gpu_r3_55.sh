#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py tests/test_fullsize_gpu.py -q -m gpu -p no:cacheprovider -x > gpurun_out/suite_tmaout.log 2>&1
echo "=== kernels + model + fullsize: exit $?"; grep -E "passed|failed|error" gpurun_out/suite_tmaout.log | tail -n 2; grep -E "^E  |^FAILED" gpurun_out/suite_tmaout.log | head -8 | cut -c1-300
run() {  # label, env...
  label=$1; shift
  env "$@" timeout 300 python bench.py --steps 20 --warmup 5 --no-extras --cpu-budget 0.2 > gpurun_out/v_$label.json 2> gpurun_out/v_$label.err || tail -2 gpurun_out/v_$label.err
  python - "$label" <<'P'
import json,sys
d=json.load(open('gpurun_out/v_%s.json' % sys.argv[1]))
L=d['launch_ms']
print('%-10s value %d step %.4f |' % (sys.argv[1], d['value'], d['ms_per_step']), ' '.join('%.4f' % v for k,v in list(L.items())[:17]))
P
}
run epw16 A=1
run prev B2I_LIB=/root/repo/ab/libb200infer_prev.so
run epw16_b A=1
run prev_b B2I_LIB=/root/repo/ab/libb200infer_prev.so
