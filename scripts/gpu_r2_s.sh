#!/bin/bash
# round 2, final evidence run: the whole GPU suite, smoke(), C2 / C4 / C3 / C1 bench lines with the final code
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/r2h_pytest_gpu.txt 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2h_pytest_gpu.txt
tail -3 gpurun_out/r2h_pytest_gpu.txt
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke OK')" > gpurun_out/r2h_smoke.txt 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2h_smoke.txt
for w in c2 c4 c3 c1; do
  timeout 900 python bench.py --workload $w --steps 3 --warmup 3 > gpurun_out/r2h_bench_$w.json 2> gpurun_out/r2h_bench_$w.err; echo "bench $w rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r2h_bench_$w.json"))
print("$w", round(d["value"],1), "GCUPS", round(d["ms_per_step"],3), "ms; e2e", round(d["e2e"]["value"],1), "; search_only", d["search_only"]["value"] and round(d["search_only"]["value"],1), "rescore ms", round(d["search_only"]["rescore_ms_per_step"],2), "topk ms", round(d["search_only"]["topk_ms_per_step"],2), "; frac", round(d["roofline"]["frac"],4), "; parity", d["parity"]["checked"], d["parity"]["mismatch"], "; cpu", round(d["cpu_baseline"]["value"],1), d["cpu_baseline"]["cores"], "; setup", d.get("setup_s"))
PY
done
