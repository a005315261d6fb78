"""A/B of the packer's stream taper inside ONE process: two packings of the same C2 database (all streams 1536 columns
vs shorter streams at the end), searches interleaved per query length, then the pipelined sweep of each.
  python scripts/ab_taper.py [--reps 5]"""
import argparse, json, os, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "seq-align-gpu_b200"))
import swgpu
from swgpu import synth

ap = argparse.ArgumentParser()
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--lqs", default=",".join(str(x) for x in synth.C2_QUERY_LENGTHS))
a = ap.parse_args()
codes, off = synth.make_db(43, 570000, lo=2, hi=35213, total=int(205e6), force_max=35213)
tab, _ = synth.load_matrix(os.path.join(ROOT, "scoring", "BLOSUM62.txt"))
def mk(cols, taper):
    os.environ["SWGPU_PACK_TAPER"] = str(taper)
    return swgpu.Database(codes, off, stream_cols=cols)
# name -> (stream_cols, halvings); the first is the baseline
variants = {"c1536t2": (1536, 2), "c1536t3": (1536, 3), "c2304t3": (2304, 3), "c3072t3": (3072, 3), "c3072t4": (3072, 4)}
dbs = {k: mk(*v) for k, v in variants.items()}
base = next(iter(dbs))
tot = {k: 0.0 for k in dbs}
for lq in [int(x) for x in a.lqs.split(",")]:
    q = synth.make_query(777, lq)
    t = {k: [] for k in dbs}
    for k, d in dbs.items():
        d.search_resident(q, tab, -10, -1)
    for r in range(a.reps):
        for k, d in dbs.items():
            t[k].append(d.search_resident(q, tab, -10, -1)["ms_total"])
    med = {k: statistics.median(v) for k, v in t.items()}
    for k in dbs:
        tot[k] += med[k]
    print(json.dumps({"lq": lq, base + "_med": round(med[base], 3)} | {k: round(med[k] / med[base], 4) for k in dbs if k != base}), flush=True)
print(json.dumps({"sum_" + k: round(v, 1) for k, v in tot.items()}))
qs = [synth.make_query(100 + i, L) for i, L in enumerate(synth.C2_QUERY_LENGTHS)]
many = {k: [] for k in dbs}
for r in range(3):
    for k, d in dbs.items():
        many[k].append(d.search_many(qs, tab, -10, -1, k=100, want_scores=False)[3]["ms_total"])
print(json.dumps({"many_" + k: [round(x, 1) for x in v] for k, v in many.items()}))
