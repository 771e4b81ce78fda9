#!/bin/bash
# one bench line at N GPUs (e2e vs device-timed check).  Usage: tools/gpu_tp_one.sh N
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
n=${1:-4}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus $n --steps 64 --warmup 5 --no-cpu-baseline --no-north-star > gpurun_out/bench_one_tp$n.json 2> gpurun_out/bench_one_tp$n.err; echo "bench tp$n exit $?"
python - $n <<'PY'
import json, sys
d = json.loads(open(f"gpurun_out/bench_one_tp{sys.argv[1]}.json").read().strip().splitlines()[-1])
print("N", d["n_gpus"], "decode", round(d["value"]), "ms", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["value"]), "e2e ms", round(d["e2e"]["ms_per_step"], 4), "clocks", d["clocks"], "parity", d["parity"]["ok"])
PY
nproc
