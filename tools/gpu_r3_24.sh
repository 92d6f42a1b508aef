#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python tools/run_job_bench.py --out gpurun_out/run_job_bench.jsonl 2>&1 | cut -c1-400
for share in 0 20; do
  B2I_JPEG_GPU_SHARE=$share B2I_TRACE=1 timeout 300 python tools/jpeg_bench.py --label "direct share $share" --reps 3 2>&1 | grep -E "arrangement|jpeg decode" | tail -2 | cut -c1-250
  B2I_JPEG_GPU_SHARE=$share timeout 300 python tools/jpeg_bench.py --label "direct share $share 3 callers" --reps 3 --callers 3 2>&1 | grep -E "arrangement" | tail -1 | cut -c1-250
done
