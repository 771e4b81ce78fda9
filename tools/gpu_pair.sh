#!/bin/bash
# GPU visit for the GEMM work: parity, microbenchmarks vs cuBLAS (pair kernel, cluster split-K), traces.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export NVC_DEBUG=1
tools/gpu_tests.sh gemm > gpurun_out/tests_gemm.log 2>&1; echo "gemm tests exit $?"; tail -5 gpurun_out/gemm.log
timeout 600 python tools/gemm_microbench.py --ks > gpurun_out/gemm_ks_micro.log 2>&1; echo "ks micro exit $?"
timeout 900 python tools/gemm_microbench.py --pair > gpurun_out/gemm_pair_micro.log 2>&1; echo "pair micro exit $?"
tools/gpu_tests.sh layers e2e > gpurun_out/tests_e2e.log 2>&1; echo "e2e tests exit $?"; cat gpurun_out/summary.txt
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_ks.json 2> gpurun_out/bench_ks.err; echo "bench exit $?"
