"""Run-to-run spread of one resident search: the same query N times in a row, per-launch K1S time and SM clock.
  python scripts/prof_repeat.py [--lqs 567,1000] [--n 30]"""
import argparse, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "seq-align-gpu_b200"))
import swgpu
from swgpu import synth

ap = argparse.ArgumentParser()
ap.add_argument("--lqs", default="567,1000")
ap.add_argument("--n", type=int, default=30)
ap.add_argument("--interleave", action="store_true", help="alternate the query lengths instead of running each N times in a row")
a = ap.parse_args()
codes, off = synth.make_db(43, 570000, lo=2, hi=35213, total=int(205e6), force_max=35213)
tab, _ = synth.load_matrix(os.path.join(ROOT, "scoring", "BLOSUM62.txt"))
db = swgpu.Database(codes, off)
lqs = [int(x) for x in a.lqs.split(",")]
qs = {lq: synth.make_query(777, lq) for lq in lqs}
def clk():
    o = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,temperature.gpu", "--format=csv,noheader,nounits", "-i", "0"],
                       capture_output=True, text=True).stdout.strip()
    return o
res = {lq: [] for lq in lqs}
order = [lq for _ in range(a.n) for lq in lqs] if a.interleave else [lq for lq in lqs for _ in range(a.n)]
for lq in order:
    st = db.search_resident(qs[lq], tab, -10, -1)
    res[lq].append(round(st["ms_inter"], 2))
for lq in lqs:
    print(json.dumps({"lq": lq, "ms_inter": res[lq], "clk_after": clk()}))
