#!/bin/bash
# TP=8: one-shot vs two-shot peer all-reduce on the 8B / 32B decode shapes
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() { local to=$1; shift; timeout "$to" python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port "$1" "${@:2}"; }
NVC_P2P_MODE=2 run 200 29561 tools/tp_check.py --model qwen3-0.6b --out gpurun_out/tp8_mode2.json > gpurun_out/tp8_mode2.log 2>&1; echo "tp8 two-shot parity exit $?"; tail -n 1 gpurun_out/tp8_mode2.log
NVC_P2P_MODE=2 run 400 29562 tools/sweep.py --model qwen3-8b --batches 64 256 --prompt-len 512 --out gpurun_out/sweep_8b_tp8_mode2.json > gpurun_out/sweep_8b_tp8_mode2.log 2>&1; echo "8b two-shot exit $?"
NVC_P2P_MODE=2 run 500 29563 tools/sweep.py --model qwen3-32b --batches 256 --prompt-len 1024 --out gpurun_out/sweep_32b_tp8_mode2.json > gpurun_out/sweep_32b_tp8_mode2.log 2>&1; echo "32b two-shot exit $?"
NVC_P2P_MODE=2 run 300 29564 bench.py --gpus 8 --steps 64 --warmup 5 > gpurun_out/bench_tp8_mode2.json 2> gpurun_out/bench_tp8_mode2.err; echo "bench two-shot exit $?"
grep -h '^{' gpurun_out/sweep_8b_tp8_mode2.log gpurun_out/sweep_32b_tp8_mode2.log gpurun_out/bench_tp8_mode2.json | cut -c1-1200
