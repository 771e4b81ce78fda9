#!/bin/bash
# Round 2, paged decode v3 (2/4 warps per CTA, K2 folded into the prologue): parity groups, warps x splits sweep,
# A/B bench lines (fused K2 on/off, forced 4-warp CTAs), full bench line.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
tools/gpu_tests.sh decode elementwise layers e2e > gpurun_out/tests_stdout.log 2>&1; echo "tests exit $?"
cat gpurun_out/summary.txt
timeout 600 python tools/decode_microbench.py --splits > gpurun_out/pd_sweep.txt 2>&1; echo "sweep exit $?"
timeout 300 python tools/decode_microbench.py > gpurun_out/pd_shapes.txt 2>&1; echo "shapes exit $?"
B="python bench.py --steps 64 --warmup 5 --no-cpu-baseline --no-north-star"
NVC_FUSED_K2=0 NVC_PD_WARPS=4 timeout 300 $B > gpurun_out/bench_unfused_w4.json 2> gpurun_out/bench_unfused_w4.err; echo "A exit $?"
NVC_FUSED_K2=0 timeout 300 $B > gpurun_out/bench_unfused_auto.json 2> gpurun_out/bench_unfused_auto.err; echo "B exit $?"
NVC_PD_WARPS=4 timeout 300 $B > gpurun_out/bench_fused_w4.json 2> gpurun_out/bench_fused_w4.err; echo "C exit $?"
timeout 300 $B > gpurun_out/bench_fused_auto.json 2> gpurun_out/bench_fused_auto.err; echo "D exit $?"
for f in unfused_w4 unfused_auto fused_w4 fused_auto; do echo $f; cut -c1-400 gpurun_out/bench_$f.json; tail -n 3 gpurun_out/bench_$f.err; done
