#!/bin/bash
# Final-state evidence of a round: launch list of the bench command (shares per kernel) and an `ncu --set full` capture of
# the paged decode kernel on exactly the launches bench.py's roofline times (NVC_NCU_RANGE + --profile-from-start off), from
# which profiles/paged_decode_ncu_traffic.json (DRAM bytes per launch at the bench shape) is written.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
BENCH="python bench.py --no-cpu-baseline --no-north-star"
timeout 200 $BENCH > gpurun_out/plain.json 2> gpurun_out/plain.err || { echo "plain bench failed"; tail -n 5 gpurun_out/plain.err; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 2000 -c 700 --csv --log-file gpurun_out/launches.csv \
    python bench.py --no-cpu-baseline --no-north-star --steps 4 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?"
NVC_NCU_RANGE=paged_decode timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:paged_decode_mma -c 3 -f -o gpurun_out/prof_pd $BENCH > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"
ncu -i gpurun_out/prof_pd.ncu-rep --page raw --csv > gpurun_out/prof_pd_raw.csv 2> /dev/null
python tools/ncu_summarize.py traffic gpurun_out/prof_pd_raw.csv paged_decode_mma gpurun_out/paged_decode_ncu_traffic.json \
    "round 2 final, bench shape (64 seqs, 8 kv heads, context 581), captured on the launches the roofline times" 64 8 581
python tools/ncu_summarize.py launches gpurun_out/launches.csv > gpurun_out/launches_summary.md 2>&1; head -n 16 gpurun_out/launches_summary.md
python tools/ncu_summarize.py full gpurun_out/prof_pd_raw.csv > gpurun_out/prof_pd_summary.md 2>&1; cat gpurun_out/prof_pd_summary.md | cut -c1-400
