#!/bin/bash
# BASELINE config 4 at TP = 2: Qwen3-14B, batch 128, 2048-token prompts (sweep: prefill tok/s, decode tok/s, fractions),
# plus the bench line at TP = 2.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() { local to=$1 port=$2; shift 2; timeout "$to" python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port "$port" "$@"; }
run 400 29571 tools/sweep.py --model qwen3-14b --batches 128 --prompt-len 2048 --out gpurun_out/sweep_14b_tp2_cfg4.json > gpurun_out/sweep_14b_tp2_cfg4.log 2>&1; echo "sweep 14b tp2 exit $?"
grep '^{' gpurun_out/sweep_14b_tp2_cfg4.log | cut -c1-900
run 300 29572 bench.py --gpus 2 --steps 64 --warmup 5 > gpurun_out/bench_tp2.json 2> gpurun_out/bench_tp2.err; echo "bench tp2 exit $?"
cut -c1-400 gpurun_out/bench_tp2.json
