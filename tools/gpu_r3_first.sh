#!/bin/bash
# Round-3 first GPU pass: whole suite, no-alloc test alone, smoke, default bench, reference arm, AlexNet bench,
# the 1000-image parity record.  Logs under gpurun_out/.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider -s > gpurun_out/gpu_suite.log 2>&1
echo "=== full -m gpu suite exit $?"; grep -E "passed|failed|error" gpurun_out/gpu_suite.log | tail -n 4 | cut -c1-300
grep -E "^E |Error" gpurun_out/gpu_suite.log | head -20 | cut -c1-300
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -p no:cacheprovider > gpurun_out/zz_alone.log 2>&1
echo "=== zz alone exit $?"; tail -n 3 gpurun_out/zz_alone.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo "=== smoke exit $?"; tail -n 2 gpurun_out/smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 --profile-out gpurun_out/launch_share.csv > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "=== bench exit $?"; tail -3 gpurun_out/bench.err; cut -c1-1500 gpurun_out/bench.json
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
echo "=== reference arm exit $?"; cut -c1-300 gpurun_out/bench_reference.json
for b in 32 256; do
  timeout 600 python bench.py --model alexnet --batch $b --steps 20 --warmup 5 --cpu-budget 5 > gpurun_out/bench_alexnet_bs$b.json 2> gpurun_out/bench_alexnet_bs$b.err
  echo "=== alexnet bs$b exit $?"; tail -2 gpurun_out/bench_alexnet_bs$b.err; cut -c1-400 gpurun_out/bench_alexnet_bs$b.json
done
timeout 1500 python tools/parity_1000.py --out gpurun_out/parity_1000.json > gpurun_out/parity_1000.log 2>&1
echo "=== parity_1000 exit $?"; tail -40 gpurun_out/parity_1000.log | cut -c1-300
