#!/bin/bash
# Multi-GPU visit (gpurun --gpus N): TP parity against the oracle, the bench line at every TP size up to N,
# and the BASELINE sweeps that fit.  Usage: tools/gpu_tp.sh N [model:tp:batches:prompt_len ...]
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-2}
nvidia-smi --query-gpu=index,name,memory.total --format=csv > gpurun_out/gpus.txt 2>&1
run() { local to=$1; shift; timeout "$to" python -m torch.distributed.run --nnodes=1 --nproc-per-node "$1" --master-addr 127.0.0.1 --master-port "$2" "${@:3}"; }
port=29520
for tp in 2 4 8; do
  [ $tp -gt $N ] && break
  port=$((port+1))
  run 200 $tp $port tools/tp_check.py --model qwen3-0.6b --out gpurun_out/tp${tp}_0.6b.json > gpurun_out/tp${tp}_0.6b.log 2>&1; echo "tp$tp 0.6b parity exit $?"
  port=$((port+1))
  run 300 $tp $port bench.py --gpus $tp --steps 64 --warmup 5 > gpurun_out/bench_tp${tp}.json 2> gpurun_out/bench_tp${tp}.err; echo "bench tp$tp exit $?"
  if [ $tp -eq $N ]; then
    port=$((port+1))
    NVC_P2P_AR=0 run 300 $tp $port bench.py --gpus $tp --steps 64 --warmup 5 > gpurun_out/bench_tp${tp}_nccl.json 2> gpurun_out/bench_tp${tp}_nccl.err; echo "bench tp$tp (NCCL all-reduce) exit $?"
  fi
done
shift
for sw in "$@"; do    # model:tp:batches:prompt_len
  IFS=: read model tp batches plen <<< "$sw"
  port=$((port+1))
  run 500 $tp $port tools/sweep.py --model $model --batches ${batches//,/ } --prompt-len $plen \
      --out gpurun_out/sweep_${model}_tp${tp}.json > gpurun_out/sweep_${model}_tp${tp}.log 2>&1; echo "sweep $sw exit $?"
done
tail -n 2 gpurun_out/tp*_0.6b.log | tail -n 30
