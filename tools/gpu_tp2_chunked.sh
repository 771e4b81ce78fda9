#!/bin/bash
# TP = 2 parity of the chunked-prefill path (mixed steps; all-reduce through NCCL) next to the
# reference scheduling mode (peer-memory all-reduce in decode), tiny and Qwen3-0.6B shapes.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() { local to=$1 port=$2; shift 2; timeout "$to" python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port "$port" "$@"; }
: > gpurun_out/tp2_chunked.jsonl
port=29560
for spec in "qwen3-tiny 24" "qwen3-tiny 0" "qwen3-0.6b 64" "qwen3-0.6b 0"; do
  set -- $spec
  port=$((port+1))
  extra=""; [ "$2" != "0" ] && extra="--chunked $2"
  run 240 $port tools/tp_check.py --model $1 $extra --out gpurun_out/tp2_$1_$2.json > gpurun_out/tp2_$1_$2.log 2>&1
  echo "tp2 $1 chunked=$2 exit $?"
  cat gpurun_out/tp2_$1_$2.json >> gpurun_out/tp2_chunked.jsonl 2>/dev/null; echo >> gpurun_out/tp2_chunked.jsonl
done
cat gpurun_out/tp2_chunked.jsonl
