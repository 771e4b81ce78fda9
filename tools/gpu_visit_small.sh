#!/bin/bash
# Small verification visit: named test groups, then the default bench line.  Usage: tools/gpu_visit_small.sh group...
cd "$(dirname "$0")/.."
tools/gpu_tests.sh "$@" > gpurun_out/tests_stdout.txt 2>&1; echo "tests exit $?"; cat gpurun_out/summary.txt
timeout 400 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench.json").read().strip().splitlines()[-1])
print("decode", round(d["value"]), "ms", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["value"]), "roofline", round(d["roofline"]["frac"], 3), "parity", d["parity"]["ok"])
print({k: v for k, v in d["breakdown"].items()})
ns = d["north_star"]; print("NS", round(ns["value"]), ns["clocks"], ns["breakdown"])
PY
