"""Profiling driver: C2-shaped database (optionally scaled), ONE query length, a few resident searches.
  python scripts/prof_one.py [--lq 375] [--scale 1.0] [--reps 3] [--topk 0]"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "seq-align-gpu_b200"))
import swgpu
from swgpu import synth

ap = argparse.ArgumentParser()
ap.add_argument("--lq", type=int, default=375)
ap.add_argument("--scale", type=float, default=1.0)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--topk", type=int, default=0)
ap.add_argument("--long-threshold", type=int, default=0)
ap.add_argument("--stream-cols", type=int, default=0)
a = ap.parse_args()
n = int(570000 * a.scale)
codes, off = synth.make_db(43, n, lo=2, hi=35213, total=int(205e6 * a.scale), force_max=35213)
tab, _ = synth.load_matrix(os.path.join(ROOT, "scoring", "BLOSUM62.txt"))
q = synth.make_query(777, a.lq)
db = swgpu.Database(codes, off, long_threshold=a.long_threshold, stream_cols=a.stream_cols)
for r in range(a.reps):
    st = db.search_resident(q, tab, -10, -1)
    if a.topk:
        db.topk(a.topk)
        st = db.last_stats()
    st["gcups"] = st["cells"] / st["ms_total"] / 1e6
    st["gcups_inter"] = st["cells_inter"] / max(st["ms_inter"], 1e-9) / 1e6
    print(json.dumps(st))
