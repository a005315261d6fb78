#!/bin/bash
rm -f gpurun_out/prof_ab.log
for lq in 128 256 512; do for c in 2 3; do
  echo "== lq $lq ctas $c" >> gpurun_out/prof_ab.log; SWGPU_K1S_CTAS=$c timeout 300 python scripts/prof_one.py --lq $lq --reps 3 >> gpurun_out/prof_ab.log 2>&1
done; done
python - <<'PY'
import json
for l in open('gpurun_out/prof_ab.log'):
    l=l.strip()
    if l.startswith('=='): print(l); continue
    try: d=json.loads(l)
    except Exception: print(l); continue
    print({k:(round(v,2) if isinstance(v,float) else v) for k,v in d.items() if k in ('ms_total','ms_intra','gcups','tile_rows','lanes_per_pair','passes')})
PY
