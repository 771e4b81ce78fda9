#!/bin/bash
# What the driver runs at round end, in its order: the GPU tests in ONE process, smoke(), the reference arm, the bench.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1100 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee gpurun_out/driver_like.txt
tail -n 5 gpurun_out/pytest_gpu.log | tee -a gpurun_out/driver_like.txt
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" | tee -a gpurun_out/driver_like.txt
tail -n 3 gpurun_out/smoke.log | tee -a gpurun_out/driver_like.txt
timeout 400 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?" | tee -a gpurun_out/driver_like.txt
cat gpurun_out/bench.json
