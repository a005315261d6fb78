"""A/B of an engine env knob inside ONE process (same GPU, interleaved runs).
  python scripts/ab_env.py --var SWGPU_K1S_CTA_ITEMS --a "" --b 1 [--lqs 144,375,1000] [--reps 5]"""
import argparse, json, os, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "seq-align-gpu_b200"))
import swgpu
from swgpu import synth

ap = argparse.ArgumentParser()
ap.add_argument("--var", required=True)
ap.add_argument("--a", default="")
ap.add_argument("--b", default="1")
ap.add_argument("--lqs", default="144,189,375,567,729,1000,2005")
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--scale", type=float, default=1.0)
a = ap.parse_args()
n = int(570000 * a.scale)
codes, off = synth.make_db(43, n, lo=2, hi=35213, total=int(205e6 * a.scale), force_max=35213)
tab, _ = synth.load_matrix(os.path.join(ROOT, "scoring", "BLOSUM62.txt"))
db = swgpu.Database(codes, off)
def setv(v):
    if v == "":
        os.environ.pop(a.var, None)
    else:
        os.environ[a.var] = v
for lq in [int(x) for x in a.lqs.split(",")]:
    q = synth.make_query(777, lq)
    db.search_resident(q, tab, -10, -1)
    ta, tb = [], []
    for r in range(a.reps):
        setv(a.a); ta.append(db.search_resident(q, tab, -10, -1)["ms_total"])
        setv(a.b); tb.append(db.search_resident(q, tab, -10, -1)["ms_total"])
    setv("")
    print(json.dumps({"lq": lq, "a_min": round(min(ta), 3), "a_med": round(statistics.median(ta), 3),
                      "b_min": round(min(tb), 3), "b_med": round(statistics.median(tb), 3),
                      "b_over_a": round(statistics.median(tb) / statistics.median(ta), 4)}))
