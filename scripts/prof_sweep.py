"""C2 sweep, serial (search_resident + topk per query) vs pipelined (swgpu_search_many).
  python scripts/prof_sweep.py [--scale 1.0] [--reps 2] [--topk 100]"""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "seq-align-gpu_b200"))
import numpy as np
import swgpu
from swgpu import synth

ap = argparse.ArgumentParser()
ap.add_argument("--scale", type=float, default=1.0)
ap.add_argument("--reps", type=int, default=2)
ap.add_argument("--topk", type=int, default=100)
ap.add_argument("--order", default="file", choices=["file", "desc"])
a = ap.parse_args()
LQS = [144, 189, 222, 375, 464, 567, 657, 729, 850, 1000, 1500, 2005, 2504, 3005, 3564, 4061, 4548, 4743, 5147, 5478]
n = int(570000 * a.scale)
codes, off = synth.make_db(43, n, lo=2, hi=35213, total=int(205e6 * a.scale), force_max=35213)
tab, _ = synth.load_matrix(os.path.join(ROOT, "scoring", "BLOSUM62.txt"))
qs = [synth.make_query(100 + i, L) for i, L in enumerate(LQS)]
if a.order == "desc":
    qs = qs[::-1]
db = swgpu.Database(codes, off)
cells = sum(len(q) for q in qs) * int(off[-1])
for r in range(a.reps):
    t0 = time.perf_counter(); dev = 0.0; ref_hits = []
    for q in qs:
        st = db.search_resident(q, tab, -10, -1)
        ref_hits.append(db.topk(a.topk) if a.topk else None)
        st = db.last_stats(); dev += st["ms_total"] + st["ms_topk"]
    t1 = time.perf_counter()
    print(json.dumps({"mode": "serial", "ms_device": dev, "ms_wall": (t1 - t0) * 1e3, "gcups_device": cells / dev / 1e6,
                      "gcups_wall": cells / (t1 - t0) / 1e9}))
    t0 = time.perf_counter()
    _, hits, per, bs = db.search_many(qs, tab, -10, -1, k=a.topk, want_scores=False)
    t1 = time.perf_counter()
    same = (hits == ref_hits) if a.topk else None
    print(json.dumps({"mode": "many", "ms_device": bs["ms_total"], "ms_wall": (t1 - t0) * 1e3,
                      "gcups_device": cells / bs["ms_total"] / 1e6, "gcups_wall": cells / (t1 - t0) / 1e9,
                      "hits_equal_serial": same, "launches": bs["kernel_launches"]}))
