#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nproc
timeout 900 python tools/jpeg_bench.py --share-sweep > gpurun_out/jpeg_share_sweep.txt 2>&1; echo "=== share sweep exit $?"
grep -v "trace" gpurun_out/jpeg_share_sweep.txt | cut -c1-330
timeout 600 python -m pytest tests/test_model_gpu.py -q -m gpu -p no:cacheprovider -x -k "jpeg or bad_image" 2>&1 | tail -3
