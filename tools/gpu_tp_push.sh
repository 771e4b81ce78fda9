#!/bin/bash
# TP visit for the GEMM + NVLS all-reduce in one kernel: kernel parity + timing (tools/nvls_check.py), e2e TP parity against the oracle
# (tools/tp_check.py), bench lines with the push on (default) and off (NVC_GEMM_AR=0).  Usage: tools/gpu_tp_push.sh N
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-2}
run() { local to=$1; shift; timeout "$to" python -m torch.distributed.run --nnodes=1 --nproc-per-node "$1" --master-addr 127.0.0.1 --master-port "$2" "${@:3}"; }
run 300 $N 29551 tools/nvls_check.py --out gpurun_out/nvls_push_tp${N}.jsonl > gpurun_out/nvls_push_tp${N}.log 2>&1; echo "nvls_check exit $?"
grep -E '"mode": 5|one_kernel|then_ll|summary' gpurun_out/nvls_push_tp${N}.log | cut -c1-220
run 200 $N 29552 tools/tp_check.py --model qwen3-0.6b --out gpurun_out/tp${N}_push.json > gpurun_out/tp${N}_push.log 2>&1; echo "tp_check exit $?"
tail -n 2 gpurun_out/tp${N}_push.log | cut -c1-300
run 300 $N 29553 bench.py --gpus $N --steps 64 --warmup 5 --no-cpu-baseline --no-north-star > gpurun_out/bench_tp${N}_push.json 2> gpurun_out/bench_tp${N}_push.err; echo "bench push exit $?"
NVC_GEMM_AR=0 run 300 $N 29554 bench.py --gpus $N --steps 64 --warmup 5 --no-cpu-baseline --no-north-star > gpurun_out/bench_tp${N}_nopush.json 2> gpurun_out/bench_tp${N}_nopush.err; echo "bench nopush exit $?"
for f in push nopush; do cut -c1-160 gpurun_out/bench_tp${N}_$f.json; tail -n 2 gpurun_out/bench_tp${N}_$f.err; done
