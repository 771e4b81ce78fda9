#!/bin/bash
# Final N = 2 check, launched as the driver launches it: TP tests, the reference arm and the full default bench line.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest -q -m gpu -p no:cacheprovider tests/test_tp_gpu.py > gpurun_out/tp_tests.log 2>&1; echo "tp tests exit $?"; tail -n 3 gpurun_out/tp_tests.log
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $T --master-port 29581 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/bench_ref_tp2.json 2> gpurun_out/bench_ref_tp2.err; echo "ref arm exit $?"; tail -c 600 gpurun_out/bench_ref_tp2.json
s=$(date +%s)
timeout 500 $T --master-port 29582 bench.py --gpus 2 > gpurun_out/bench_tp2_default.json 2> gpurun_out/bench_tp2_default.err; echo "bench tp2 exit $? in $(( $(date +%s) - s )) s"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_tp2_default.json").read().strip().splitlines()[-1])
ns = d.get("north_star", {})
print("N", d["n_gpus"], "decode", round(d["value"]), "e2e", round(d["e2e"]["value"]), "clocks", d["clocks"], "parity", d["parity"]["ok"], d["parity"].get("ranks_agree"),
      "roofline", round(d["roofline"]["frac"], 3), "NS", round(ns.get("value", 0)), ns.get("parallelism"))
PY
