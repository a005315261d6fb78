"""C4 (35 213-aa query x 1 024 targets up to 35 213 aa): timing + parity of a sample against the oracle.
  python scripts/prof_c4.py [--reps 2] [--check 6]"""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "seq-align-gpu_b200"))
import numpy as np
import swgpu
from swgpu import synth

ap = argparse.ArgumentParser()
ap.add_argument("--reps", type=int, default=2)
ap.add_argument("--check", type=int, default=0, help="targets checked against the scalar oracle (slow: 1.2 G cells each)")
ap.add_argument("--lq", type=int, default=35213)
a = ap.parse_args()
q, codes, off = synth.config4(lq=a.lq)
tab, _ = synth.load_matrix(os.path.join(ROOT, "scoring", "BLOSUM62.txt"))
db = swgpu.Database(codes, off)
for r in range(a.reps):
    st = db.search_resident(q, tab, -10, -1)
    st["gcups"] = st["cells"] / st["ms_total"] / 1e6
    print(json.dumps({k: (round(v, 3) if isinstance(v, float) else v) for k, v in st.items()}))
if a.check:
    from oracle import oracle as O
    s = db.scores()
    idx = [0, 1, 15, 16, 63, 64, 500][: a.check]
    for i in idx:
        want = O.score32(q, codes[off[i]:off[i + 1]], tab, -10, -1)
        print("target", i, "len", int(off[i + 1] - off[i]), "gpu", int(s[i]), "oracle", int(want), "OK" if s[i] == want else "MISMATCH")
