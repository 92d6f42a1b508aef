#!/bin/bash
# Two GPUs: in-process sharder tests, bench.py in one process (inproc record), torchrun N=2 both arms.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python -m pytest tests/test_multigpu_gpu.py -q -m gpu -p no:cacheprovider 2>&1 | tail -3
timeout 600 python bench.py --steps 20 --warmup 5 --cpu-budget 2 > gpurun_out/bench_inproc2.json 2> gpurun_out/bench_inproc2.err; echo "=== bench (1 process, 2 visible) exit $?"; tail -2 gpurun_out/bench_inproc2.err
python -c "
import json; d=json.load(open('gpurun_out/bench_inproc2.json')); print('value',round(d['value']),'e2e',round(d['e2e']['value'])); print(json.dumps(d['inproc'])[:900])"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; echo "=== torchrun N=2 exit $?"; tail -2 gpurun_out/bench_2gpu.err | cut -c1-300
python -c "
import json; d=json.loads([l for l in open('gpurun_out/bench_2gpu.json') if l.startswith('{')][-1]); print('N=2 value',round(d['value']),'e2e',round(d['e2e']['value']), 'ceiling', d['e2e'].get('h2d_ceiling'), 'sustained', d['sustained'], 'cpu', d['cpu_baseline'])"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_2gpu_ref.json 2> gpurun_out/bench_2gpu_ref.err; echo "=== torchrun reference N=2 exit $?"; cut -c1-200 gpurun_out/bench_2gpu_ref.json
