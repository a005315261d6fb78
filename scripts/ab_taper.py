"""A/B of the packer's stream taper inside ONE process: two packings of the same C2 database (all streams 1536 columns
vs shorter streams at the end), searches interleaved per query length, then the pipelined sweep of each.
  python scripts/ab_taper.py [--reps 5]"""
import argparse, json, os, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "seq-align-gpu_b200"))
import swgpu
from swgpu import synth

ap = argparse.ArgumentParser()
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--lqs", default=",".join(str(x) for x in synth.C2_QUERY_LENGTHS))
a = ap.parse_args()
codes, off = synth.make_db(43, 570000, lo=2, hi=35213, total=int(205e6), force_max=35213)
tab, _ = synth.load_matrix(os.path.join(ROOT, "scoring", "BLOSUM62.txt"))
dbs = {"flat": swgpu.Database(codes, off, no_taper=1), "taper": swgpu.Database(codes, off)}
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
    print(json.dumps({"lq": lq, "flat_med": round(med["flat"], 3), "taper_med": round(med["taper"], 3),
                      "flat_min": round(min(t["flat"]), 3), "taper_min": round(min(t["taper"]), 3),
                      "taper_over_flat": round(med["taper"] / med["flat"], 4)}), flush=True)
print(json.dumps({"sum_flat": round(tot["flat"], 1), "sum_taper": round(tot["taper"], 1)}))
qs = [synth.make_query(100 + i, L) for i, L in enumerate(synth.C2_QUERY_LENGTHS)]
many = {k: [] for k in dbs}
for r in range(3):
    for k, d in dbs.items():
        many[k].append(d.search_many(qs, tab, -10, -1, k=100, want_scores=False)[3]["ms_total"])
print(json.dumps({"many_flat": [round(x, 1) for x in many["flat"]], "many_taper": [round(x, 1) for x in many["taper"]]}))
