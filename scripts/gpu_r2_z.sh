#!/bin/bash
# round 2, batch z: K1S slot pointer per plan: every plan against the oracle, parity suites, C2 / C4 lines
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_plans.py tests/test_gpu_parity.py tests/test_gpu_c2.py tests/test_gpu_many.py -m gpu -x -q > gpurun_out/r2z_pytest.txt 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2z_pytest.txt
timeout 900 python bench.py --workload c2 --steps 3 --warmup 3 > gpurun_out/r2z_bench_c2.json 2> gpurun_out/r2z_bench_c2.err; echo "bench c2 rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/r2z_bench_c2.json"))
print("c2", round(d["value"],1), "GCUPS", round(d["ms_per_step"],3), "ms; e2e", round(d["e2e"]["value"],1), "; frac", round(d["roofline"]["frac"],4), "; parity", d["parity"]["checked"], d["parity"]["mismatch"], d["topk_crc"])
PY
timeout 600 python bench.py --workload c4 --steps 3 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('c4', round(d['value'],1), round(d['ms_per_step'],2), 'k1', round(d['roofline']['k1_ms_per_step'],2), d['parity']['mismatch'])"
