#!/bin/bash
# round 2, batch z2: best (nested loop, slot pointer) combination per rows-per-lane value: every plan against the oracle, C2 line
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_plans.py tests/test_gpu_planner.py tests/test_gpu_c2.py -m gpu -x -q > gpurun_out/r2z2_pytest.txt 2>&1; echo "pytest rc=$?"; tail -1 gpurun_out/r2z2_pytest.txt
timeout 900 python bench.py --workload c2 --steps 3 --warmup 3 > gpurun_out/r2z2_bench_c2.json 2> gpurun_out/r2z2_bench_c2.err; echo "bench c2 rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/r2z2_bench_c2.json"))
print("c2", round(d["value"],1), "GCUPS", round(d["ms_per_step"],3), "ms; e2e", round(d["e2e"]["value"],1), "; frac", round(d["roofline"]["frac"],4), "; parity", d["parity"]["checked"], d["parity"]["mismatch"], d["topk_crc"])
PY
