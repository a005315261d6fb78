#!/bin/bash
# round 2, batch b: new GPU tests (positive gaps, SWGPU_GPUS) + bench lines for c1 / c3 / c4, then memcheck on the small case
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_positive_gaps.py tests/test_gpu_multi.py tests/test_boundary.py -m gpu -x -q > gpurun_out/r2_pytest_b.log 2>&1; echo "rc=$?" >> gpurun_out/r2_pytest_b.log
tail -n 12 gpurun_out/r2_pytest_b.log
for w in c1 c3 c4; do
  timeout 600 python bench.py --workload $w --steps 5 --warmup 3 > gpurun_out/r2_bench_$w.json 2> gpurun_out/r2_bench_$w.err; echo "bench $w rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r2_bench_$w.json"))
print("$w", round(d["value"],1), "GCUPS", round(d["ms_per_step"],3), "ms; e2e", round(d["e2e"]["value"],1), "; cpu", d["cpu_baseline"] and round(d["cpu_baseline"]["value"],2), "; parity", d["parity"]["checked"], d["parity"]["mismatch"], "; rescored", d["rescored_per_step"])
PY
done
bash scripts/gpu_sanitize.sh memcheck
