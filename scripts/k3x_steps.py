"""How long is ONE column step of the 32-bit wavefront kernel (k3x, csrc/sw_k3x.cuh), and what does a stripe hand-over cost?
A database of `--targets` equal targets of `--lt` residues with long_threshold = 1 sends every target through the 32-bit path
(targets >= 4096 residues take k3x); the query length sets the number of 128-row stripes.  Prints one JSON line per query length:
  ms_rescore, stripes, cycles per column step if the stripes were perfectly pipelined (ms * clock / lt), and the same after
  taking out a lag of `lag` columns per stripe fitted from the one-stripe and many-stripe runs.
    python scripts/k3x_steps.py [--lt 35213] [--targets 1] [--lqs 128,256,...]"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "seq-align-gpu_b200"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--lt", type=int, default=35213)
    ap.add_argument("--targets", type=int, default=1)
    ap.add_argument("--lqs", default="128,256,512,1024,4096,16384,35213")
    ap.add_argument("--reps", type=int, default=4)
    ap.add_argument("--mhz", type=float, default=1965.0)
    a = ap.parse_args()
    import swgpu
    from swgpu import synth
    g = synth.rng(5)
    lens = np.full(a.targets, a.lt, dtype=np.int64)
    off = np.zeros(a.targets + 1, dtype=np.int64); np.cumsum(lens, out=off[1:])
    codes = synth.residues(g, off[-1])
    tab = synth.load_matrix(os.path.join(ROOT, "scoring", "BLOSUM62.txt"))[0]
    db = swgpu.Database(codes, off, long_threshold=1)
    rows = []
    for lq in [int(x) for x in a.lqs.split(",")]:
        q = synth.make_query(7, lq)
        best = None
        for _ in range(a.reps):
            _, st = db.search(q, tab, -10, -1)
            ms = st["ms_rescore"]
            best = ms if best is None else min(best, ms)
        ns = (lq + 127) // 128
        rows.append(dict(lq=lq, stripes=ns, targets=a.targets, lt=a.lt, ms_rescore=round(best, 4),
                         cycles_per_col=round(best * a.mhz * 1e3 / a.lt, 1),
                         gcups=round(lq * a.lt * a.targets / best / 1e6, 1)))
        print(json.dumps(rows[-1]), flush=True)
    if len(rows) >= 2 and rows[0]["stripes"] == 1:
        c0 = rows[0]["ms_rescore"] * a.mhz * 1e3 / a.lt
        for r in rows[1:]:
            extra = r["ms_rescore"] * a.mhz * 1e3 - c0 * a.lt
            print(json.dumps(dict(lq=r["lq"], stripes=r["stripes"], one_stripe_cycles_per_col=round(c0, 1),
                                  extra_cycles_per_added_stripe=round(extra / (r["stripes"] - 1), 0),
                                  as_columns_of_lag=round(extra / (r["stripes"] - 1) / c0, 1))), flush=True)
    db.close()


if __name__ == "__main__":
    main()
