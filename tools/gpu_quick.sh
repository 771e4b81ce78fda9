#!/bin/bash
# Short visit after a kernel change: the parity groups named on the command line (default: decode e2e), then the bench line.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/diag.log
t() { local n=$1 to=$2; shift 2; timeout "$to" python -m pytest -q -m gpu -p no:cacheprovider "$@" > "gpurun_out/$n.log" 2>&1; echo "$n exit $? : $(tail -n 1 gpurun_out/$n.log)"; }
for g in ${@:-decode e2e}; do
  case $g in
    decode) t decode 240 tests/test_gemm_attn_gpu.py -k paged_decode ;;
    prefill) t prefill 240 tests/test_gemm_attn_gpu.py -k prefill ;;
    gemm) t gemm 300 tests/test_gemm_attn_gpu.py -k gemm ;;
    e2e) t e2e 400 tests/test_e2e_gpu.py ;;
    layers) t layers 200 tests/test_layers_gpu.py ;;
    elementwise) t elementwise 240 tests/test_kernels_gpu.py ;;
    pmicro) timeout 200 python tools/prefill_microbench.py gpurun_out/prefill_microbench.json > gpurun_out/prefill_microbench.log 2>&1; echo "pmicro exit $?"; tail -n 8 gpurun_out/prefill_microbench.log ;;
    micro) timeout 200 python tools/decode_microbench.py > gpurun_out/decode_microbench.log 2>&1; echo "micro exit $?"; tail -n 12 gpurun_out/decode_microbench.log ;;
    sweep8) timeout 300 python tools/sweep.py --model qwen3-8b --batches 1 64 256 --out gpurun_out/sweep_8b.json > gpurun_out/sweep_8b.log 2>&1; echo "sweep8 exit $?"; grep -o '"batch": [0-9]*\|"decode_tok_s": [0-9]*\|"paged_attn_frac": [0-9.]*' gpurun_out/sweep_8b.log | paste - - - ;;
  esac
done
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench exit $?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_quick.json"))
print("value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "roofline", d["roofline"]["frac"], d["roofline"]["us_per_launch"], d["breakdown"])
PY
