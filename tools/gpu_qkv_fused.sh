#!/bin/bash
# K3 + K2 fused (QKV GEMM epilogue): parity groups + prefill A/B through bench.py (NVC_QKV_FUSED=1|0)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python -m pytest -q -m gpu -p no:cacheprovider tests/test_gemm_attn_gpu.py -k "qkv_proj or gemm_bf16" 2>&1 | tail -8
timeout 200 python -m pytest -q -m gpu -p no:cacheprovider tests/test_kernels_gpu.py -k "qk_norm or rope" 2>&1 | tail -3
for f in 1 0; do
  NVC_QKV_FUSED=$f timeout 300 python bench.py --steps 16 --warmup 3 --no-cpu-baseline > gpurun_out/bench_qkvfused$f.json 2> gpurun_out/bench_qkvfused$f.err
  python - $f <<'PY'
import json, sys
d = json.loads(open(f"gpurun_out/bench_qkvfused{sys.argv[1]}.json").read().strip().splitlines()[-1])
print("fused", sys.argv[1], "0.6B prefill", round(d["prefill"]["tok_per_s"]), "8B prefill", round(d["north_star"]["prefill"]["tok_per_s"]),
      "parity", d["parity"]["ok"], "ns clock", d["north_star"]["clocks"]["sm_mhz"])
PY
done
