#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for cta in 0 46 100 146; do
  B2I_PAIR_TRACE=1 B2I_PAIR_TRACE_CTA=$cta timeout 120 python tools/profile_step.py 1 1 > gpurun_out/trace_$cta.log 2>&1
  echo "=== CTA $cta"
  grep "pair trace" gpurun_out/trace_$cta.log | tail -6 | cut -c1-330
done
