#!/bin/bash
# round 2, batch: k3_intra with the shifted cell form (one-operation row-to-row chain) and placement by level: tests that reach it, C3 / C1 / C2 lines
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_plans.py tests/test_gpu_upstream.py tests/test_boundary.py tests/test_host_layer.py tests/test_positive_gaps.py tests/test_gpu_many.py tests/test_gpu_c5.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2y_pytest_k3.txt 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2y_pytest_k3.txt
timeout 200 python scripts/sanitize_case.py > gpurun_out/r2y_sanitize_case_plain.txt 2>&1; echo "sanitize_case rc=$?"
for w in c3 c1; do
  timeout 900 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('$w', round(d['value'],1), round(d['ms_per_step'],2), 'search', round(d['search_only']['ms_per_step'],2), 'rescore', round(d['search_only']['rescore_ms_per_step'],2), 'topk', round(d['search_only']['topk_ms_per_step'],2), d['parity']['mismatch'])"
done
