#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for pool in 0 1; do for share in 10 20; do
  echo "--- NO_POOL=$pool share=$share"
  B2I_JPEG_NO_POOL=$pool B2I_JPEG_GPU_SHARE=$share B2I_TRACE=1 timeout 300 python tools/jpeg_bench.py --label "nopool $pool share $share" --reps 2 2>&1 | grep -E "arrangement|jpeg decode" | tail -3 | cut -c1-250
done; done
