#!/bin/bash
rm -f gpurun_out/prof_ab.log
run() { echo "== lq $1 plan $2" >> gpurun_out/prof_ab.log; SWGPU_K1S_PLAN=$2 timeout 300 python scripts/prof_one.py --lq $1 --reps 3 >> gpurun_out/prof_ab.log 2>&1; }
run 1000 32,32; run 1000 32,16; run 2005 32,32; run 2005 32,16; run 1280 32,20; run 1280 32,28; run 512 32,16; run 512 16,32; run 375 32,12; run 375 16,24
python - <<'PY'
import json
for l in open('gpurun_out/prof_ab.log'):
    l=l.strip()
    if l.startswith('=='): print(l); continue
    try: d=json.loads(l)
    except Exception: print(l); continue
    print({k:(round(v,2) if isinstance(v,float) else v) for k,v in d.items() if k in ('ms_total','ms_intra','gcups','tile_rows','lanes_per_pair','passes')})
PY
