#!/bin/bash
# round 2, probe k: k3x items stripe-major inside a generation of targets: milestone trace (profiling build), step probe, tests that reach k3x, C4 / C3 / C2 lines
mkdir -p gpurun_out; rm -f gpurun_out/r2k_k3x_trace.txt
timeout 300 python scripts/k3x_steps.py > gpurun_out/r2k_k3x_steps.jsonl 2> gpurun_out/r2k_k3x_steps.err; echo "k3x_steps rc=$?"
timeout 300 python scripts/k3x_steps.py --targets 64 --lqs 128,1024,4096,35213 --reps 2 > gpurun_out/r2k_k3x_steps_64.jsonl 2>> gpurun_out/r2k_k3x_steps.err
cat gpurun_out/r2k_k3x_steps.jsonl gpurun_out/r2k_k3x_steps_64.jsonl | grep ms_rescore
timeout 1500 python -m pytest tests/test_gpu_topk.py tests/test_gpu_traceback.py tests/test_gpu_upstream.py tests/test_gpu_parity.py tests/test_gpu_many.py tests/test_gpu_c5.py -m gpu -x -q > gpurun_out/r2_pytest_k.log 2>&1; echo "rc=$?" >> gpurun_out/r2_pytest_k.log
tail -4 gpurun_out/r2_pytest_k.log
for w in c4 c3; do
  timeout 900 python bench.py --workload $w --steps 3 --warmup 3 > gpurun_out/r2k_bench_$w.json 2> gpurun_out/r2k_bench_$w.err; echo "bench $w rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r2k_bench_$w.json"))
print("$w", round(d["value"],1), "GCUPS", round(d["ms_per_step"],3), "ms; search_only", d["search_only"]["value"] and round(d["search_only"]["value"],1), round(d["search_only"]["ms_per_step"],2), "rescore ms", round(d["search_only"]["rescore_ms_per_step"],2), "topk ms", round(d["search_only"]["topk_ms_per_step"],2), "; frac", round(d["roofline"]["frac"],4), "; parity", d["parity"]["checked"], d["parity"]["mismatch"])
PY
done
