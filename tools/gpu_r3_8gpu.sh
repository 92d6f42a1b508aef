#!/bin/bash
# Eight GPUs of one box: in-process sharder tests, bench.py in one process (inproc + dual-job record), torchrun N=2,4,8
# for both arms, batch sweep over all GPUs and the dual job (tools/sweep.py).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi -L | wc -l; nproc
timeout 300 python -m pytest tests/test_multigpu_gpu.py -q -m gpu -p no:cacheprovider 2>&1 | tail -2
timeout 400 python bench.py --steps 20 --warmup 5 --cpu-budget 0.3 > gpurun_out/bench_inproc8.json 2> gpurun_out/bench_inproc8.err; echo "=== bench (1 process, 8 visible) exit $?"; tail -2 gpurun_out/bench_inproc8.err
python -c "
import json; d=json.load(open('gpurun_out/bench_inproc8.json')); print('value',round(d['value']),'e2e',round(d['e2e']['value'])); print(json.dumps(d['inproc'])[:1200])"
port=29520
for n in 2 4 8; do
  port=$((port+1))
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/bench_${n}gpu.json 2> gpurun_out/bench_${n}gpu.err; echo "=== torchrun N=$n exit $?"; tail -1 gpurun_out/bench_${n}gpu.err | cut -c1-300
  python -c "
import json; d=json.loads([l for l in open('gpurun_out/bench_${n}gpu.json') if l.startswith('{')][-1]); print('N=$n value',round(d['value']),'e2e',round(d['e2e']['value']), 'ceiling GB/s', round(d['e2e']['h2d_ceiling']['gbs_aggregate'],1), 'bound', round(d['e2e']['h2d_ceiling']['images_per_s_bound']), 'sustained', round(d['sustained']['images_per_s']), d['sustained']['sm_mhz_median'])"
  port=$((port+1))
  [ $n = 8 ] && timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --impl reference --gpus $n --steps 4 --warmup 3 > gpurun_out/bench_${n}gpu_ref.json 2> gpurun_out/bench_${n}gpu_ref.err; echo "=== reference N=$n exit $?"; cut -c1-160 gpurun_out/bench_${n}gpu_ref.json
done
SWEEP_ONLY_ALL_GPUS=1 timeout 420 python tools/sweep.py gpurun_out/sweep_8gpu.jsonl both > gpurun_out/sweep8.log 2>&1; echo "=== sweep exit $?"; tail -3 gpurun_out/sweep8.log; grep '"kind": "dual"' gpurun_out/sweep_8gpu.jsonl | cut -c1-600
