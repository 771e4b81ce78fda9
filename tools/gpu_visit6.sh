#!/bin/bash
# Round 2: hoisted fused-K2 prologue, split-K partials summed by the norm (decode, TP 1), lm_head through the pair kernel.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
tools/gpu_tests.sh gemm decode layers e2e > gpurun_out/tests_stdout.log 2>&1; echo "tests exit $?"
cat gpurun_out/summary.txt
B="python bench.py --steps 64 --warmup 5 --no-cpu-baseline --no-north-star"
NVC_SPLITK_NORM=0 NVC_FUSED_K2=0 timeout 300 $B > gpurun_out/bench_A_plain.json 2> gpurun_out/bench_A.err; echo "A exit $?"
NVC_SPLITK_NORM=0 timeout 300 $B > gpurun_out/bench_B_fused.json 2> gpurun_out/bench_B.err; echo "B exit $?"
NVC_FUSED_K2=0 timeout 300 $B > gpurun_out/bench_C_partial.json 2> gpurun_out/bench_C.err; echo "C exit $?"
timeout 300 $B > gpurun_out/bench_D_both.json 2> gpurun_out/bench_D.err; echo "D exit $?"
for f in A_plain B_fused C_partial D_both; do echo $f; cut -c1-200 gpurun_out/bench_$f.json; done
timeout 400 python tools/lm_head_sweep.py > gpurun_out/lm_head_sweep.txt 2>&1; echo "lm_head exit $?"
cat gpurun_out/lm_head_sweep.txt
