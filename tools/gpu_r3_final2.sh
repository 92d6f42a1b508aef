#!/bin/bash
# Round-3 closing one-GPU pass: suite, no-alloc test alone, smoke, bench lines (default / fp16 / reference arm / AlexNet), ncu launch list.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider > gpurun_out/gpu_suite.log 2>&1
echo "=== full -m gpu suite exit $?"; grep -E "passed|failed|error" gpurun_out/gpu_suite.log | tail -n 3 | cut -c1-300
grep -E "^E  |^FAILED" gpurun_out/gpu_suite.log | head -20 | cut -c1-400
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -p no:cacheprovider > gpurun_out/zz_alone.log 2>&1; echo "=== zz alone exit $?"; tail -n 2 gpurun_out/zz_alone.log | cut -c1-200
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "=== smoke exit $?"; tail -n 1 gpurun_out/smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 --profile-out gpurun_out/launch_share.csv > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "=== bench exit $?"; tail -3 gpurun_out/bench.err; python - <<'P'
import json
d=json.load(open('gpurun_out/bench.json'))
print('value',round(d['value']),'e2e',round(d['e2e']['value']),'ms',round(d['ms_per_step'],4),'sustained',d['sustained'],'\nroofline',{k:d['roofline'][k] for k in ('achieved','frac','frac_burst','frac_sustained','traffic','whole_step','whole_step_sustained_segment')},'\nfp16',d['fp16'],'parity',d['parity'], '\nfrom_files', d['from_files'], '\ncpu', d['cpu_baseline'], '\n', d['launch_ms'])
P
timeout 600 python bench.py --precision fp16 --steps 20 --warmup 5 --cpu-budget 1 > gpurun_out/bench_fp16.json 2> gpurun_out/bench_fp16.err; echo "=== fp16 bench exit $?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "=== reference arm exit $?"; cut -c1-200 gpurun_out/bench_reference.json
for b in 32 256; do timeout 600 python bench.py --model alexnet --batch $b --steps 20 --warmup 5 --cpu-budget 5 > gpurun_out/bench_alexnet_bs$b.json 2> gpurun_out/bench_alexnet_bs$b.err; echo "=== alexnet bs$b exit $?"; python -c "
import json; d=json.load(open('gpurun_out/bench_alexnet_bs$b.json')); print(round(d['value']), round(d['e2e']['value']), d['ms_per_step'], d['roofline']['achieved'], d['sustained'])"; done
python tools/profile_step.py 2 1 > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 40 -c 20 --csv --log-file gpurun_out/ncu_launches.csv python tools/profile_step.py 2 1 > gpurun_out/ncu_l.log 2>&1
echo "=== ncu launch list exit $?"
