#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_topk.py tests/test_host_layer.py -m gpu -x -q > gpurun_out/pytest_topk.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_topk.log
tail -n 8 gpurun_out/pytest_topk.log
rm -f gpurun_out/prof_topk.log
for lq in 375 5478; do timeout 300 python scripts/prof_one.py --lq $lq --reps 3 --topk 100 >> gpurun_out/prof_topk.log 2>&1; done
python - <<'PY'
import json
for l in open('gpurun_out/prof_topk.log'):
    try: d=json.loads(l)
    except Exception: print(l.strip()); continue
    print({k:(round(v,2) if isinstance(v,float) else v) for k,v in d.items() if k in ('ms_total','ms_topk','topk_launches','gcups')})
PY
