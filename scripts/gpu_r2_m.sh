#!/bin/bash
# A/B: CTAs per SM of the reversed coordinate pass (SWGPU_K3X_REV_CTAS) on C4 and C2
for c in 4 2 1; do
  echo "== rev ctas $c"
  SWGPU_K3X_REV_CTAS=$c timeout 600 python bench.py --workload c4 --steps 3 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('c4', round(d['value'],1), round(d['ms_per_step'],2), 'topk', round(d['search_only']['topk_ms_per_step'],2), d['parity']['mismatch'], d['topk_crc'])"
done
