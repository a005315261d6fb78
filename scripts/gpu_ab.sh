#!/bin/bash
set -x
mkdir -p gpurun_out; rm -f gpurun_out/prof_ab.log
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 15 gpurun_out/pytest_gpu.log
for lq in 144 375 1000 2005 5478; do
  echo "== lq $lq k1s" >> gpurun_out/prof_ab.log; timeout 300 python scripts/prof_one.py --lq $lq --reps 3 >> gpurun_out/prof_ab.log 2>&1
done
for lq in 375; do
  echo "== lq $lq chunk" >> gpurun_out/prof_ab.log; timeout 300 python scripts/prof_one.py --lq $lq --reps 3 --chunk >> gpurun_out/prof_ab.log 2>&1
  for sc in 512 1024 3072; do echo "== lq $lq k1s stream_cols $sc" >> gpurun_out/prof_ab.log; timeout 300 python scripts/prof_one.py --lq $lq --reps 3 --stream-cols $sc >> gpurun_out/prof_ab.log 2>&1; done
done
python - <<'PY'
import json
for l in open('gpurun_out/prof_ab.log'):
    l=l.strip()
    if l.startswith('=='): print(l); continue
    try: d=json.loads(l)
    except Exception: print(l); continue
    print({k:(round(v,2) if isinstance(v,float) else v) for k,v in d.items() if k in ('ms_total','ms_inter','ms_intra','ms_rescore','gcups','gcups_inter','tile_rows','lanes_per_pair','passes','n_rescored')})
PY
