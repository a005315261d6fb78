"""Time every admissible K1S plan (lanes per pair g, rows per lane R) for each query length of the C2 sweep.
  python scripts/tune_plan.py [--scale 1.0]"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "seq-align-gpu_b200"))
import swgpu
from swgpu import synth

ap = argparse.ArgumentParser()
ap.add_argument("--scale", type=float, default=1.0)
ap.add_argument("--lqs", default="")
a = ap.parse_args()
LQS = [int(x) for x in a.lqs.split(",")] if a.lqs else list(synth.C2_QUERY_LENGTHS) + [64, 100, 128, 256, 300, 512, 768, 1024]
n = int(570000 * a.scale)
codes, off = synth.make_db(43, n, lo=2, hi=35213, total=int(205e6 * a.scale), force_max=35213)
tab, _ = synth.load_matrix(os.path.join(ROOT, "scoring", "BLOSUM62.txt"))
db = swgpu.Database(codes, off)
MULTI_R = (16, 20, 24, 25, 26, 27, 28, 29, 30, 31, 32)
for lq in sorted(LQS):
    q = synth.make_query(777, lq)
    os.environ.pop("SWGPU_K1S_PLAN", None)
    best = None
    res = {}
    st = db.search_resident(q, tab, -10, -1); st = db.search_resident(q, tab, -10, -1)
    default = (st["lanes_per_pair"], st["tile_rows"], st["passes"], st["ms_total"])
    plans = set()
    for g in (8, 16, 32):
        R = max(-(-lq // g), 8)
        for d in (0, 1, 2):
            if R + d <= 32:
                plans.add((g, R + d))
    for R in MULTI_R:
        if 32 * R < lq or True:
            plans.add((32, R))
    for g, R in sorted(plans):
        np_ = -(-lq // (g * R))
        if np_ > 1 and not (g == 32 and R in MULTI_R):
            continue
        os.environ["SWGPU_K1S_PLAN"] = f"{g},{R}"
        ms = 1e9
        for _ in range(2):
            st = db.search_resident(q, tab, -10, -1)
            ms = min(ms, st["ms_total"])
        if (st["lanes_per_pair"], st["tile_rows"]) != (g, R):
            continue
        res[f"{g},{R}x{st['passes']}"] = round(ms, 3)
    os.environ.pop("SWGPU_K1S_PLAN", None)
    b = min(res, key=res.get)
    print(json.dumps({"lq": lq, "default": f"{default[0]},{default[1]}x{default[2]}", "default_ms": round(default[3], 3),
                      "best": b, "best_ms": res[b], "gain": round(default[3] / res[b], 3), "all": res}))
