#!/bin/bash
# The allocation-free entry point's test alone in a fresh process, then the driver-like sequence (all GPU tests in one process,
# smoke, bench).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python -m pytest -q -m gpu -p no:cacheprovider tests/test_zz_no_alloc.py > gpurun_out/no_alloc_alone.log 2>&1; echo "alone exit $?"
tail -n 15 gpurun_out/no_alloc_alone.log
tools/gpu_driver_like.sh
