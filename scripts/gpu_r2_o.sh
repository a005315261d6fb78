#!/bin/bash
# round 2, A/B o: K1S latency form of the cell (one-operation row-to-row chain, V unshifted) in the multi-stripe kernels:
# parity of every plan with the variant, C4 (all streams are X items), and the C2 per-length A/B
mkdir -p gpurun_out
SWGPU_LIB=seq-align-gpu_b200/libswgpu_vlat.so timeout 900 python -m pytest tests/test_gpu_plans.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r2o_pytest_lat.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2o_pytest_lat.log
for rep in 1 2; do for v in "" _vlat; do
  echo "== libswgpu$v.so"
  SWGPU_LIB=seq-align-gpu_b200/libswgpu$v.so timeout 600 python bench.py --workload c4 --steps 3 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('c4', round(d['value'],1), round(d['ms_per_step'],2), 'search', round(d['search_only']['ms_per_step'],2), 'k1', round(d['roofline']['k1_ms_per_step'],2), 'rescore', round(d['search_only']['rescore_ms_per_step'],2), 'topk', round(d['search_only']['topk_ms_per_step'],2), d['parity']['mismatch'])"
done; done
timeout 1200 python scripts/ab_lib.py --a seq-align-gpu_b200/libswgpu.so --b seq-align-gpu_b200/libswgpu_vlat.so --rounds 1 > gpurun_out/r2o_ab_lat.log 2>&1; echo "ab rc=$?"; grep -v '"lib"' gpurun_out/r2o_ab_lat.log | cut -c1-220
