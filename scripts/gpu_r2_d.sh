#!/bin/bash
# round 2, batch d: K3X early stop in the coordinate passes: top-k / coordinates tests, then C4 and C2 bench lines
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_topk.py tests/test_gpu_traceback.py tests/test_gpu_upstream.py tests/test_gpu_c2.py tests/test_gpu_c5.py tests/test_gpu_c5_device.py tests/test_gpu_many.py tests/test_gpu_multi.py -m gpu -x -q -s > gpurun_out/r2_pytest_d.log 2>&1; echo "rc=$?" >> gpurun_out/r2_pytest_d.log
grep -E "coordinates,|passed|failed|rc=" gpurun_out/r2_pytest_d.log | tail -5
for w in c4 c2; do
  timeout 900 python bench.py --workload $w --steps 3 --warmup 3 > gpurun_out/r2d_bench_$w.json 2> gpurun_out/r2d_bench_$w.err; echo "bench $w rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r2d_bench_$w.json"))
print("$w", round(d["value"],1), "GCUPS", round(d["ms_per_step"],3), "ms; search_only", d["search_only"]["value"] and round(d["search_only"]["value"],1), round(d["search_only"]["ms_per_step"],2), "topk ms", round(d["search_only"]["topk_ms_per_step"],2), "; frac", round(d["roofline"]["frac"],4), "; parity", d["parity"]["checked"], d["parity"]["mismatch"])
PY
done
