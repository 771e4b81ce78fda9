#!/bin/bash
# Runs the GPU parity tests group by group, each in its own process with a timeout, so that a
# faulting kernel (sticky CUDA error) or a hang in one group cannot take the others down.
# Logs go to gpurun_out/ (merged back by gpurun).  Usage: tools/gpu_tests.sh [group ...]
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/diag.log
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
rc=0
: > gpurun_out/summary.txt
run() {  # name, timeout, pytest args...
  local name=$1 to=$2; shift 2
  echo "=== $name" | tee -a gpurun_out/summary.txt
  timeout "$to" python -m pytest -q -m gpu -p no:cacheprovider "$@" > "gpurun_out/$name.log" 2>&1
  local r=$?
  tail -n 4 "gpurun_out/$name.log" | tee -a gpurun_out/summary.txt
  echo "exit $r" | tee -a gpurun_out/summary.txt
  [ $r -ne 0 ] && rc=1
}
groups=${@:-elementwise gemm decode prefill guard layers e2e loader tp}
for g in $groups; do
  case $g in
    elementwise) run elementwise 300 tests/test_kernels_gpu.py ;;
    gemm) run gemm 400 tests/test_gemm_attn_gpu.py -k gemm ;;
    decode) run decode 300 tests/test_gemm_attn_gpu.py -k paged_decode ;;
    prefill) run prefill 300 tests/test_gemm_attn_gpu.py -k prefill ;;
    guard) run guard 300 tests/test_guard_bands_gpu.py ;;
    e2e) run e2e 900 tests/test_e2e_gpu.py ;;
    layers) run layers 300 tests/test_layers_gpu.py ;;
    loader) run loader 300 tests/test_hf_loader.py ;;
    tp) run tp 700 tests/test_tp_gpu.py ;;
    *) run "$(basename "$g" .py)" 900 "$g" ;;
  esac
done
echo "=== diag"; tail -n 60 gpurun_out/diag.log
exit $rc
