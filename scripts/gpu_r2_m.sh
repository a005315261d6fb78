#!/bin/bash
# round 2, A/B m: k3x CTAs per SM: C4 line (top-100 coordinates on long targets), 64-target throughput case, C2 top-k
mkdir -p gpurun_out
for v in "" _vctas1 _vctas2 _vctas3; do
  echo "== libswgpu$v.so"
  SWGPU_LIB=seq-align-gpu_b200/libswgpu$v.so timeout 300 python scripts/k3x_steps.py --targets 64 --lqs 35213 --reps 2 2>/dev/null | grep ms_rescore | cut -c1-140
  SWGPU_LIB=seq-align-gpu_b200/libswgpu$v.so timeout 600 python bench.py --workload c4 --steps 3 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('c4', round(d['value'],1), round(d['ms_per_step'],2), 'search', round(d['search_only']['ms_per_step'],2), 'rescore', round(d['search_only']['rescore_ms_per_step'],2), 'topk', round(d['search_only']['topk_ms_per_step'],2), d['parity']['mismatch'])"
done
