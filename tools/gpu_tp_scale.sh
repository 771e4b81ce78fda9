#!/bin/bash
# Scaling visit on an 8-GPU box: e2e TP parity against the oracle at TP 8, then the bench line at N = 8, 4, 2 (the driver's
# SCALE run does the same back to back).  Usage: tools/gpu_tp_scale.sh
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() { local to=$1; shift; timeout "$to" python -m torch.distributed.run --nnodes=1 --nproc-per-node "$1" --master-addr 127.0.0.1 --master-port "$2" "${@:3}"; }
run 200 8 29561 tools/tp_check.py --model qwen3-0.6b --out gpurun_out/tp8_check.json > gpurun_out/tp8_check.log 2>&1; echo "tp_check 8 exit $?"
tail -n 1 gpurun_out/tp8_check.log | cut -c1-300
port=29562
for n in 8 4 2; do
  port=$((port+1))
  run 300 $n $port bench.py --gpus $n --steps 64 --warmup 5 --no-cpu-baseline --no-north-star > gpurun_out/bench_scale_tp$n.json 2> gpurun_out/bench_scale_tp$n.err; echo "bench tp$n exit $?"
  python - $n <<'PY'
import json, sys
try:
    d = json.loads(open(f"gpurun_out/bench_scale_tp{sys.argv[1]}.json").read().strip().splitlines()[-1])
    print("N", d["n_gpus"], "decode", round(d["value"]), "ms", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["value"]), "prefill", round(d["prefill"]["tok_per_s"]),
          "parity", d["parity"]["ok"], d["parity"].get("ranks_agree"), {k: v for k, v in d["breakdown"].items() if k in ("paged_decode_us", "allreduce_us", "gemm_o_us", "gemm_down_us")})
except Exception as e:
    print("no line", e)
PY
done
