#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 60 python tools/lm_head_one.py || exit 1
timeout 100 ncu --set full --clock-control none --import-source on -k regex:gemm_pair_kernel -c 4 -f -o gpurun_out/prof_lm_head python tools/lm_head_one.py > gpurun_out/ncu_lm_head.log 2>&1
echo "ncu exit $?"
ncu -i gpurun_out/prof_lm_head.ncu-rep --page raw --csv > gpurun_out/prof_lm_head_raw.csv 2> /dev/null
python tools/ncu_summarize.py full gpurun_out/prof_lm_head_raw.csv > gpurun_out/prof_lm_head_summary.md 2>&1; cut -c1-600 gpurun_out/prof_lm_head_summary.md
