#!/bin/bash
# N = 2: the TP parity tests, then one short bench line.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest -q -m gpu -p no:cacheprovider tests/test_tp_gpu.py > gpurun_out/tp_tests.log 2>&1; echo "tp tests exit $?"; tail -n 3 gpurun_out/tp_tests.log
tools/gpu_tp_one.sh 2
