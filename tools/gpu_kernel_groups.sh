#!/bin/bash
# Runs the GPU kernel-parity tests one group per process (a trapped kernel poisons its CUDA context,
# so groups are isolated) and leaves one log per group under gpurun_out/.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu_info.csv 2>&1
rc=0
for k in "gather_cpasyncB" "gather_tmaB" "im2col" "smallc" "linear" "fp32_mode" "maxpool or avgpool or normalize or softmax or resize"; do
  name=$(echo "$k" | tr ' ' '_' | cut -c1-24)
  timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "$k" -p no:cacheprovider > "gpurun_out/kern_${name}.log" 2>&1
  r=$?
  echo "=== group [$k] exit $r"
  grep -E "^(FAILED|ERROR|[0-9]+ (passed|failed))|passed|failed" "gpurun_out/kern_${name}.log" | tail -25
  grep -E "AssertionError: |B2IError" "gpurun_out/kern_${name}.log" | cut -c1-400 | sort | uniq -c | sort -rn | head -12
  [ $r -ne 0 ] && rc=1
done
exit $rc
