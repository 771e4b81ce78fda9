#!/bin/bash
# One GPU-box visit sized for a short budget: new-feature tests first, then the bench line, the ncu
# launch list and one `--set full` capture of the paged decode kernel, then the rest of the suite.
# Usage: tools/gpu_visit.sh [new] [bench] [ncu] [rest]
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
what=${@:-new bench ncu rest}
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
t() { local n=$1 to=$2; shift 2; timeout "$to" python -m pytest -q -m gpu -p no:cacheprovider "$@" > "gpurun_out/$n.log" 2>&1; echo "$n exit $? : $(tail -n 1 gpurun_out/$n.log)"; }
for w in $what; do
  case $w in
    new)
      t prefill 240 tests/test_gemm_attn_gpu.py -k prefill
      t sampling 120 tests/test_kernels_gpu.py -k sample
      t e2e 400 tests/test_e2e_gpu.py
      t loader 200 tests/test_hf_loader.py ;;
    bench)
      timeout 400 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; cat gpurun_out/bench.json ;;
    ncu)
      timeout 200 $BENCH > gpurun_out/plain.log 2>&1 &&
      timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 600 --csv \
          --log-file gpurun_out/launches.csv $BENCH > gpurun_out/ncu_launches.log 2>&1
      echo "ncu launches exit $?"
      timeout 600 ncu --set full --clock-control none --import-source on \
          -k regex:'paged_decode_mma' -s 30 -c 3 -f -o gpurun_out/prof_pd $BENCH > gpurun_out/ncu_full.log 2>&1
      echo "ncu full exit $?"
      ncu -i gpurun_out/prof_pd.ncu-rep --page raw --csv > gpurun_out/prof_pd_raw.csv 2> /dev/null ;;
    rest)
      t elementwise 300 tests/test_kernels_gpu.py -k "not sample"
      t gemm 400 tests/test_gemm_attn_gpu.py -k gemm
      t decode 300 tests/test_gemm_attn_gpu.py -k paged_decode
      t layers 300 tests/test_layers_gpu.py ;;
  esac
done
tail -n 40 gpurun_out/diag.log 2> /dev/null
