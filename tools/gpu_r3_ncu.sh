#!/bin/bash
# One `ncu --set full` capture of the 20 launches of a ResNet-18 step (after the same command has run without ncu).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python tools/profile_step.py 2 1 > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none -s 40 -c 20 -f -o gpurun_out/prof_step python tools/profile_step.py 2 1 > gpurun_out/ncu_f.log 2>&1
echo "=== ncu full exit $?"
ncu -i gpurun_out/prof_step.ncu-rep --page raw --csv > gpurun_out/ncu_step_raw.csv 2>/dev/null; rm -f gpurun_out/prof_step.ncu-rep; ls -la gpurun_out/ncu_step_raw.csv
