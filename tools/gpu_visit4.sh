#!/bin/bash
# Final-state evidence: full bench line (CPU legs included), ncu launch list of the bench command, ncu --set full
# capture of the paged decode kernel (DRAM traffic per launch -> profiles/paged_decode_ncu_traffic.json).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
timeout 400 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; cut -c1-200 gpurun_out/bench.json
timeout 120 $BENCH > gpurun_out/plain.log 2>&1 &&
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 600 --csv \
    --log-file gpurun_out/launches.csv $BENCH > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:'paged_decode_mma' -s 30 -c 3 -f \
    -o gpurun_out/prof_pd $BENCH > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"
ncu -i gpurun_out/prof_pd.ncu-rep --page raw --csv > gpurun_out/prof_pd_raw.csv 2> /dev/null
