#!/bin/bash
# ncu --set full with source counters for launches 1 (layer1 halo-patch) and 5 (layer2.0.conv1+downsample) of a step.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python tools/profile_step.py 2 1 > gpurun_out/plain.log 2>&1 && \
ncu --set full --import-source on --clock-control none -s 41 -c 5 -f -o gpurun_out/prof_l1_l20 python tools/profile_step.py 2 1 > gpurun_out/ncu_f.log 2>&1
echo "=== ncu exit $?"; ls -la gpurun_out/prof_l1_l20.ncu-rep
