#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
B2I_LIB=/root/repo/ab/libb200infer_trace.so B2I_PATCH_TRACE=1 timeout 120 python tools/profile_step.py 1 1 > gpurun_out/trace_patch.log 2>&1
grep -A24 "patch trace" gpurun_out/trace_patch.log | tail -52 | cut -c1-160
