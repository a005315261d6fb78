#!/bin/bash
# round 2: C5 at a tenth of the named size (device-generated UniRef50-shaped database, 64 queries, BLOSUM45, top-100) with the final code
mkdir -p gpurun_out
timeout 1200 python bench.py --workload c5 --scale 0.1 --steps 1 --warmup 1 > gpurun_out/r2h_bench_c5_tenth.json 2> gpurun_out/r2h_bench_c5_tenth.err; echo "rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/r2h_bench_c5_tenth.json"))
print("c5/10", round(d["value"],1), "GCUPS", round(d["ms_per_step"],1), "ms; e2e", round(d["e2e"]["value"],1), "; frac", d["roofline"]["frac"] and round(d["roofline"]["frac"],4), "; parity", d.get("parity"))
PY
