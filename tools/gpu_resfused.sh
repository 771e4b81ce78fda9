#!/bin/bash
# residual add in the row-parallel GEMM epilogue (prefill): parity + prefill A/B through bench.py (NVC_RESIDUAL_FUSED=1|0)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python -m pytest -q -m gpu -p no:cacheprovider tests/test_gemm_attn_gpu.py -k "residual_add or qkv_proj" 2>&1 | tail -6
timeout 300 python -m pytest -q -m gpu -p no:cacheprovider tests/test_kernels_gpu.py tests/test_e2e_gpu.py tests/test_layers_gpu.py tests/test_guard_bands_gpu.py 2>&1 | tail -3
for f in 1 0; do
  NVC_RESIDUAL_FUSED=$f timeout 300 python bench.py --steps 16 --warmup 3 --no-cpu-baseline > gpurun_out/bench_resfused$f.json 2> gpurun_out/bench_resfused$f.err
  python - $f <<'PY'
import json, sys
d = json.loads(open(f"gpurun_out/bench_resfused{sys.argv[1]}.json").read().strip().splitlines()[-1])
ns = d["north_star"]
print("fused", sys.argv[1], "0.6B prefill", round(d["prefill"]["tok_per_s"]), "8B prefill", round(ns["prefill"]["tok_per_s"]),
      "parity", d["parity"]["ok"], "ns consistency", ns.get("consistency", {}).get("ok"), "ns clock", ns["clocks"]["sm_mhz"])
PY
done
