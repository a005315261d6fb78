#!/bin/bash
# round 2, A/B w: every multi-stripe plan (rows per lane) with and without the nested column loop, two query lengths
mkdir -p gpurun_out
for v in "" _vnested; do
  SWGPU_LIB=seq-align-gpu_b200/libswgpu$v.so timeout 900 python scripts/tune_plan.py --lqs 2005,4743 > gpurun_out/r2w_tune$v.jsonl 2>&1; echo "rc=$?"
done
python - <<'PY'
import json
a=[json.loads(l) for l in open('gpurun_out/r2w_tune.jsonl') if l.startswith('{')]
b=[json.loads(l) for l in open('gpurun_out/r2w_tune_vnested.jsonl') if l.startswith('{')]
for x,y in zip(a,b):
    print('lq',x['lq'])
    for k in x['all']:
        if k in y['all']: print('  ',k, x['all'][k], y['all'][k], round(y['all'][k]/x['all'][k],4))
PY
