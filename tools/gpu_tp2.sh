#!/bin/bash
# Two-shot peer all-reduce check: parity + bench with NVC_P2P_MODE=1 (one-shot) and =2 (two-shot) at TP = $1
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-2}
run() { local to=$1; shift; timeout "$to" python -m torch.distributed.run --nnodes=1 --nproc-per-node "$1" --master-addr 127.0.0.1 --master-port "$2" "${@:3}"; }
port=29540
for mode in 2 1; do
  port=$((port+1))
  NVC_P2P_MODE=$mode run 200 $N $port tools/tp_check.py --model qwen3-0.6b --out gpurun_out/tp${N}_mode${mode}.json > gpurun_out/tp${N}_mode${mode}.log 2>&1; echo "tp$N mode $mode parity exit $?"
  tail -n 1 gpurun_out/tp${N}_mode${mode}.log
  port=$((port+1))
  NVC_P2P_MODE=$mode run 300 $N $port bench.py --gpus $N --steps 64 --warmup 5 > gpurun_out/bench_tp${N}_mode${mode}.json 2> gpurun_out/bench_tp${N}_mode${mode}.err; echo "bench tp$N mode $mode exit $?"
  for sw in "${@:2}"; do
    IFS=: read model batches plen <<< "$sw"
    port=$((port+1))
    NVC_P2P_MODE=$mode run 500 $N $port tools/sweep.py --model $model --batches ${batches//,/ } --prompt-len $plen \
        --out gpurun_out/sweep_${model}_tp${N}_mode${mode}.json > gpurun_out/sweep_${model}_tp${N}_mode${mode}.log 2>&1; echo "sweep $sw mode $mode exit $?"
  done
done
