#!/bin/bash
# GPU-box visit for the chunked-prefill milestone, most valuable first (a visit may be cut short):
# e2e parity (mixed steps included), BASELINE config 3 job in both scheduling modes, the bench line,
# the 0.6B batch sweep, the ncu launch list, then the remaining parity groups.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/diag.log
t() { local n=$1 to=$2; shift 2; timeout "$to" python -m pytest -q -m gpu -p no:cacheprovider "$@" > "gpurun_out/$n.log" 2>&1; echo "$n exit $? : $(tail -n 1 gpurun_out/$n.log)"; }
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
t e2e 400 tests/test_e2e_gpu.py
timeout 400 python tools/config3.py --budget 2048 8192 --out gpurun_out/config3.json > gpurun_out/config3.log 2>&1; echo "config3 exit $?"
timeout 400 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; cat gpurun_out/bench.json
timeout 200 python tools/sweep.py --model qwen3-0.6b --batches 1 64 256 --out gpurun_out/sweep_06b.json > gpurun_out/sweep_06b.log 2>&1; echo "sweep 0.6b exit $?"
timeout 120 $BENCH > gpurun_out/plain.log 2>&1 &&
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 600 --csv \
    --log-file gpurun_out/launches.csv $BENCH > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?"
t layers 200 tests/test_layers_gpu.py
t loader 200 tests/test_hf_loader.py
t prefill 240 tests/test_gemm_attn_gpu.py -k prefill
t decode 240 tests/test_gemm_attn_gpu.py -k paged_decode
t gemm 300 tests/test_gemm_attn_gpu.py -k gemm
t elementwise 240 tests/test_kernels_gpu.py
tail -n 60 gpurun_out/diag.log 2> /dev/null
