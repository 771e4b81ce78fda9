#!/bin/bash
# decode GEMMs with a 64-row A sub-tile: parity under both ring budgets, bench A/B (NVC_GEMM_TRIM=0|1|2)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for t in 1 2; do
  NVC_GEMM_TRIM=$t timeout 300 python -m pytest -q -m gpu -p no:cacheprovider tests/test_gemm_attn_gpu.py -k "gemm" 2>&1 | tail -2
done
timeout 200 python -m pytest -q -m gpu -p no:cacheprovider tests/test_e2e_gpu.py tests/test_layers_gpu.py 2>&1 | tail -2
for t in 0 1 2 0 1; do
  NVC_GEMM_TRIM=$t timeout 300 python bench.py --steps 64 --warmup 5 --no-cpu-baseline --no-north-star > gpurun_out/bench_trim$t.json 2> gpurun_out/bench_trim$t.err
  python - $t <<'PY'
import json, sys
d = json.loads(open(f"gpurun_out/bench_trim{sys.argv[1]}.json").read().strip().splitlines()[-1])
print("trim", sys.argv[1], "decode", round(d["value"]), "ms", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["value"]), "parity", d["parity"]["ok"],
      {k: v for k, v in d["breakdown"].items() if k.startswith("gemm")})
PY
done
