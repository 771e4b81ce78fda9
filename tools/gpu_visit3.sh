#!/bin/bash
# Whole-model check after the prefill attention rewrite: e2e parity, bench, 0.6B / 8B sweeps, config 3 job,
# ncu --set full capture of the new prefill kernel on the 32B/TP8 shape (8 x 4096 tokens).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/diag.log
t() { local n=$1 to=$2; shift 2; timeout "$to" python -m pytest -q -m gpu -p no:cacheprovider "$@" > "gpurun_out/$n.log" 2>&1; echo "$n exit $? : $(tail -n 1 gpurun_out/$n.log)"; }
t e2e 400 tests/test_e2e_gpu.py
t layers 200 tests/test_layers_gpu.py
timeout 400 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; cat gpurun_out/bench.json
timeout 200 python tools/sweep.py --model qwen3-0.6b --batches 1 64 256 --out gpurun_out/sweep_06b.json > gpurun_out/sweep_06b.log 2>&1; echo "sweep 0.6b exit $?"
timeout 300 python tools/sweep.py --model qwen3-8b --batches 1 64 256 --out gpurun_out/sweep_8b.json > gpurun_out/sweep_8b.log 2>&1; echo "sweep 8b exit $?"
timeout 400 python tools/config3.py --budget 2048 --out gpurun_out/config3.json > gpurun_out/config3.log 2>&1; echo "config3 exit $?"
PREFILL_SHAPE=4 timeout 300 ncu --set full --clock-control none --import-source on -k regex:prefill_attn_v2 -s 6 -c 2 -f \
    -o gpurun_out/prof_prefill_v2 python tools/prefill_microbench.py > gpurun_out/ncu_prefill.log 2>&1; echo "ncu prefill exit $?"
ncu -i gpurun_out/prof_prefill_v2.ncu-rep --page raw --csv > gpurun_out/prof_prefill_v2_raw.csv 2> /dev/null
tail -n 30 gpurun_out/diag.log 2> /dev/null
