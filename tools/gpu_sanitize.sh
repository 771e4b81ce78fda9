#!/bin/bash
# compute-sanitizer over the GPU kernel tests (SURVEY 5 stance): memcheck on the elementwise / decode / GEMM / prefill
# parity tests, racecheck (shared-memory hazards) on the kernels that hand data between warps through shared memory
# outside of mbarrier-ordered TMA traffic.  Summaries -> gpurun_out/sanitize_*.txt
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
run() {  # name, tool, timeout, pytest args
  local name=$1 tool=$2 to=$3; shift 3
  echo "=== $name ($tool)"
  timeout "$to" $CS --tool "$tool" --launch-timeout 0 --print-limit 20 --log-file "gpurun_out/sanitize_$name.log" \
      python -m pytest -q -m gpu -x -p no:cacheprovider "$@" > "gpurun_out/sanitize_$name.pytest.log" 2>&1
  echo "exit $?"
  tail -n 3 "gpurun_out/sanitize_$name.pytest.log"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY" "gpurun_out/sanitize_$name.log" | tail -n 2
}
run mem_elementwise memcheck 600 tests/test_kernels_gpu.py
run mem_decode memcheck 600 tests/test_gemm_attn_gpu.py -k "paged_decode"
run mem_gemm memcheck 900 tests/test_gemm_attn_gpu.py -k "gemm_partial or gemm_silu or gemm_reference or gemm_strided"
run mem_prefill memcheck 600 tests/test_gemm_attn_gpu.py -k "prefill_attn and not paged"
run race_elementwise racecheck 600 tests/test_kernels_gpu.py -k "rmsnorm or qk_norm or argmax or sample"
run race_decode racecheck 600 tests/test_gemm_attn_gpu.py -k "paged_decode_fused or paged_decode_attn"
for f in gpurun_out/sanitize_*.log; do case $f in *pytest.log) ;; *) echo "--- $f"; grep -vE "^=+ *$" "$f" | head -n 40;; esac; done > gpurun_out/sanitize_summary.txt
