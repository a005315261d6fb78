#!/bin/bash
# round 2, A/B n: K1S X-item wait as a loop the whole warp leaves together
mkdir -p gpurun_out
for rep in 1 2; do for v in "" _vuwait; do
  echo "== libswgpu$v.so"
  SWGPU_LIB=seq-align-gpu_b200/libswgpu$v.so timeout 600 python bench.py --workload c4 --steps 3 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('c4', round(d['value'],1), round(d['ms_per_step'],2), 'search', round(d['search_only']['ms_per_step'],2), 'k1', round(d['roofline']['k1_ms_per_step'],2), 'rescore', round(d['search_only']['rescore_ms_per_step'],2), 'topk', round(d['search_only']['topk_ms_per_step'],2), d['parity']['mismatch'])"
done; done
