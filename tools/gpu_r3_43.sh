#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
B2I_LIB=/root/repo/ab/libb200infer_trace.so B2I_PAIR_TRACE=1 timeout 120 python tools/profile_step.py 1 1 > gpurun_out/trace_0.log 2>&1
grep -A9 "pair trace\] CTA" gpurun_out/trace_0.log | tail -22 | cut -c1-330
