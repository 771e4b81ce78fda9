#!/bin/bash
# One GPU-box visit: parity tests, the bench line, the reference arm, the ncu launch list + one
# `--set full` capture of the top kernels, the 8B sweep and the kernel microbenchmarks.
# Everything lands in gpurun_out/.  Usage: tools/gpu_round.sh [tests] [bench] [ncu] [sweep] [micro]
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
what=${@:-tests bench ncu sweep micro}
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
for w in $what; do
  case $w in
    tests) tools/gpu_tests.sh elementwise gemm decode prefill layers e2e > gpurun_out/tests.log 2>&1; echo "tests exit $?" ;;
    bench)
      timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; cat gpurun_out/bench.json
      timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref exit $?" ;;
    ncu)
      timeout 300 $BENCH > gpurun_out/plain.log 2>&1 &&
      timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 600 --csv \
          --log-file gpurun_out/launches.csv $BENCH > gpurun_out/ncu_launches.log 2>&1
      echo "ncu launches exit $?"
      timeout 900 ncu --set full --clock-control none --import-source on \
          -k regex:'paged_decode|gemm_bf16' -s 40 -c 8 -f -o gpurun_out/prof $BENCH > gpurun_out/ncu_full.log 2>&1
      echo "ncu full exit $?" ;;
    sweep)
      timeout 900 python tools/sweep.py --model qwen3-8b --batches 1 64 256 --prompt-len 512 --out gpurun_out/sweep_8b.json \
          > gpurun_out/sweep_8b.log 2>&1; echo "sweep exit $?"; cat gpurun_out/sweep_8b.log | tail -5 ;;
    micro)
      timeout 600 python tools/decode_microbench.py > gpurun_out/decode_micro.log 2>&1; echo "decode micro exit $?"
      timeout 600 python tools/gemm_microbench.py > gpurun_out/gemm_micro.log 2>&1; echo "gemm micro exit $?" ;;
  esac
done
