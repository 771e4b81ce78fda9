#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python -m pytest -q -m gpu -p no:cacheprovider -x tests/test_gemm_attn_gpu.py -k "lm_head_swapped" 2>&1 | tail -n 5
: > gpurun_out/lm_head_sweep_v.txt
for v in 75968 37984 18992; do NVC_SWAP_MIN_N=100000000 timeout 200 python tools/lm_head_sweep.py $v 2>&1 | grep -E "K=1024|M=64 K=4096" | tee -a gpurun_out/lm_head_sweep_v.txt; done
timeout 400 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench.json").read().strip().splitlines()[-1])
print("decode", round(d["value"]), "ms", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["value"]), "roofline", round(d["roofline"]["frac"], 3), "parity", d["parity"])
print({k: v for k, v in d["breakdown"].items()})
ns = d["north_star"]; print("NS", round(ns["value"]), ns["clocks"], ns["breakdown"])
PY
