#!/bin/bash
# quick GPU check: parity suite + a few single-query timings
mkdir -p gpurun_out; rm -f gpurun_out/prof_ab.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 6 gpurun_out/pytest_gpu.log
for lq in ${LQS:-144 375 1000 5478}; do
  echo "== lq $lq" >> gpurun_out/prof_ab.log; timeout 300 python scripts/prof_one.py --lq $lq --reps 3 $EXTRA >> gpurun_out/prof_ab.log 2>&1
done
python - <<'PY'
import json
for l in open('gpurun_out/prof_ab.log'):
    l=l.strip()
    if l.startswith('=='): print(l); continue
    try: d=json.loads(l)
    except Exception: print(l); continue
    print({k:(round(v,2) if isinstance(v,float) else v) for k,v in d.items() if k in ('ms_total','ms_inter','ms_rescore','gcups','tile_rows','lanes_per_pair','passes','n_rescored')})
PY
