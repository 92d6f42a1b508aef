#!/bin/bash
# ncu --set full of the stem (both forms) and of the CTA-pair launches of one step.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export B2I_NPROD=3
python tools/profile_step.py 2 1 > gpurun_out/plain1.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:stem_kernel -s 2 -c 1 -f -o gpurun_out/prof_stem1 python tools/profile_step.py 2 1 > gpurun_out/ncu_stem1.log 2>&1
echo "stem1 ncu exit $?"; tail -3 gpurun_out/ncu_stem1.log
B2I_STEM=2 python tools/profile_step.py 2 1 > gpurun_out/plain2.log 2>&1 && \
B2I_STEM=2 ncu --set full --clock-control none --import-source on -k regex:stem2_kernel -s 2 -c 1 -f -o gpurun_out/prof_stem2 python tools/profile_step.py 2 1 > gpurun_out/ncu_stem2.log 2>&1
echo "stem2 ncu exit $?"; tail -3 gpurun_out/ncu_stem2.log
python tools/profile_step.py 2 1 > gpurun_out/plain3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:conv_umma_pair_kernel -s 18 -c 9 -f -o gpurun_out/prof_pair python tools/profile_step.py 2 1 > gpurun_out/ncu_pair.log 2>&1
echo "pair ncu exit $?"; tail -3 gpurun_out/ncu_pair.log
ls -la gpurun_out/*.ncu-rep
