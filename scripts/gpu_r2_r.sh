#!/bin/bash
# round 2, batch r: forward coordinate pass bounded to the rows that can hold the hit's score: coordinate tests, C4 / C2 lines, A/B knob
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_topk.py tests/test_gpu_traceback.py tests/test_gpu_upstream.py tests/test_gpu_parity.py tests/test_gpu_many.py tests/test_gpu_c5.py tests/test_gpu_c5_device.py tests/test_gpu_plans.py -m gpu -x -q > gpurun_out/r2_pytest_r.log 2>&1; echo "rc=$?" >> gpurun_out/r2_pytest_r.log
tail -4 gpurun_out/r2_pytest_r.log
for knob in 0 1; do
  if [ $knob = 1 ]; then export SWGPU_TOPK_NO_ROW_BOUND=1; echo "== without the row bound"; fi
  timeout 900 python bench.py --workload c4 --steps 3 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('c4', round(d['value'],1), round(d['ms_per_step'],2), 'search', round(d['search_only']['ms_per_step'],2), 'rescore', round(d['search_only']['rescore_ms_per_step'],2), 'topk', round(d['search_only']['topk_ms_per_step'],2), d['parity']['mismatch'], d['topk_crc'])"
done
unset SWGPU_TOPK_NO_ROW_BOUND
timeout 900 python bench.py --workload c2 --steps 3 --warmup 3 > gpurun_out/r2r_bench_c2.json 2> gpurun_out/r2r_bench_c2.err; echo "bench c2 rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/r2r_bench_c2.json"))
print("c2", round(d["value"],1), "GCUPS", round(d["ms_per_step"],3), "ms; topk ms", round(d["search_only"]["topk_ms_per_step"],2), "; frac", round(d["roofline"]["frac"],4), "; parity", d["parity"]["checked"], d["parity"]["mismatch"], d["topk_crc"])
PY
