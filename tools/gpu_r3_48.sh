#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
B2I_LIB=/root/repo/ab/libb200infer_trace.so B2I_STEM2_TRACE=1 timeout 120 python tools/profile_step.py 1 1 > gpurun_out/trace_stem2.log 2>&1
grep -A40 "stem2 trace" gpurun_out/trace_stem2.log | tail -42 | cut -c1-200
