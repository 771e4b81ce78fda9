#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python -m pytest -q -m gpu -p no:cacheprovider tests/test_kernels_gpu.py -k "argmax or sample" 2>&1 | tail -n 25
timeout 200 python tools/head_microbench.py 2>&1 | tee gpurun_out/head_microbench.txt
