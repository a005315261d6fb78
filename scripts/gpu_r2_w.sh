#!/bin/bash
# round 2, A/B w2: every multi-stripe plan under the four combinations of (nested column loop, slot pointer in registers)
mkdir -p gpurun_out
for v in n0s0 n0s1 n1s0 n1s1; do
  SWGPU_LIB=seq-align-gpu_b200/libswgpu_v$v.so timeout 900 python scripts/tune_plan.py --lqs 2005 > gpurun_out/r2w2_tune_$v.jsonl 2>&1; echo "$v rc=$?"
done
SWGPU_LIB=seq-align-gpu_b200/libswgpu.so timeout 900 python scripts/tune_plan.py --lqs 2005 > gpurun_out/r2w2_tune_intree.jsonl 2>&1
python - <<'PY'
import json
R={}
for v in ("n0s0","n0s1","n1s0","n1s1","intree"):
    R[v]=[json.loads(l) for l in open(f'gpurun_out/r2w2_tune_{v}.jsonl') if l.startswith('{')][0]['all']
for k in R["n0s0"]:
    b=R["n0s0"][k]
    print(f"{k:10s} n0s0 {b:8.3f}  " + "  ".join(f"{v} {R[v][k]/b:.4f}" for v in ("n0s1","n1s0","n1s1","intree")))
PY
