#!/bin/bash
# Round-3 eleventh pass: suite, JPEG test repeated, TMEM-A chain microbenchmark, default bench lines, ncu tables of one step.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/ -q -m gpu -p no:cacheprovider > gpurun_out/gpu_suite.log 2>&1
echo "=== full -m gpu suite exit $?"; grep -E "passed|failed|error" gpurun_out/gpu_suite.log | tail -n 3 | cut -c1-300
grep -E "^E  |^FAILED" gpurun_out/gpu_suite.log | head -20 | cut -c1-300
for i in 1 2 3 4 5; do timeout 300 python -m pytest tests/test_model_gpu.py -q -m gpu -p no:cacheprovider -k "jpeg or bad_image" 2>&1 | tail -1; done
timeout 300 tools/microbench/umma_rate > gpurun_out/umma_rate.log 2>&1; echo "=== umma_rate exit $?"; grep -E "TMEM|cta_group 1" gpurun_out/umma_rate.log
timeout 900 python bench.py --steps 20 --warmup 5 --profile-out gpurun_out/launch_share.csv > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "=== bench exit $?"; tail -3 gpurun_out/bench.err; python - <<'P'
import json
d=json.load(open('gpurun_out/bench.json'))
print('value',round(d['value']),'e2e',round(d['e2e']['value']),'ms',round(d['ms_per_step'],4),'sustained',d['sustained'],'\nroofline',{k:d['roofline'][k] for k in ('achieved','frac','frac_burst','frac_sustained','peak_used','whole_step','whole_step_sustained_segment')},'\nfp16',d['fp16'],'parity',d['parity'])
P
timeout 600 python bench.py --precision fp16 --steps 20 --warmup 5 --cpu-budget 1 > gpurun_out/bench_fp16.json 2> gpurun_out/bench_fp16.err; echo "=== fp16 bench exit $?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "=== reference arm exit $?"; cut -c1-200 gpurun_out/bench_reference.json
for b in 32 256; do timeout 600 python bench.py --model alexnet --batch $b --steps 20 --warmup 5 --cpu-budget 5 > gpurun_out/bench_alexnet_bs$b.json 2> gpurun_out/bench_alexnet_bs$b.err; echo "=== alexnet bs$b exit $?"; python -c "
import json; d=json.load(open('gpurun_out/bench_alexnet_bs$b.json')); print(round(d['value']), round(d['e2e']['value']), d['ms_per_step'], d['roofline']['achieved'], d['sustained'])"; done
python tools/profile_step.py 2 1 > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 40 -c 20 --csv --log-file gpurun_out/ncu_launches.csv python tools/profile_step.py 2 1 > gpurun_out/ncu_l.log 2>&1
echo "=== ncu launch list exit $?"
python tools/profile_step.py 2 1 > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none -s 40 -c 20 -f -o gpurun_out/prof_step python tools/profile_step.py 2 1 > gpurun_out/ncu_f.log 2>&1
echo "=== ncu full exit $?"
ncu -i gpurun_out/prof_step.ncu-rep --page raw --csv > gpurun_out/ncu_step_raw.csv 2>/dev/null; ls -la gpurun_out/prof_step.ncu-rep gpurun_out/ncu_step_raw.csv
