"""A/B of two builds of libswgpu.so on ONE box: child processes alternate A, B, A, B ... (SWGPU_LIB),
each timing every C2 query length (min of --reps search_resident calls) and one pipelined sweep.
  python scripts/ab_lib.py --a seq-align-gpu_b200/libswgpu.so --b seq-align-gpu_b200/libswgpu_b.so [--rounds 2]"""
import argparse, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LQS = [144, 189, 222, 375, 464, 567, 657, 729, 850, 1000, 1500, 2005, 2504, 3005, 3564, 4061, 4548, 4743, 5147, 5478]

def child(reps):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "seq-align-gpu_b200"))
    import swgpu
    from swgpu import synth
    codes, off = synth.make_db(43, 570000, lo=2, hi=35213, total=int(205e6), force_max=35213)
    tab, _ = synth.load_matrix(os.path.join(ROOT, "scoring", "BLOSUM62.txt"))
    db = swgpu.Database(codes, off)
    qs = [synth.make_query(100 + i, L) for i, L in enumerate(LQS)]
    out = {"lib": swgpu.LIB_PATH, "ms": {}}
    for q in qs:
        db.search_resident(q, tab, -10, -1)
        out["ms"][len(q)] = round(min(db.search_resident(q, tab, -10, -1)["ms_total"] for _ in range(reps)), 3)
    import zlib
    out["crc"] = {len(q): zlib.crc32(db.search(q, tab, -10, -1)[0].tobytes()) for q in (qs[0], qs[3], qs[9], qs[11])}
    db.search_many(qs, tab, -10, -1, k=100, want_scores=False)
    out["many_ms"] = round(min(db.search_many(qs, tab, -10, -1, k=100, want_scores=False)[3]["ms_total"] for _ in range(2)), 2)
    out["many_gcups"] = round(sum(LQS) * int(off[-1]) / out["many_ms"] / 1e6, 1)
    print(json.dumps(out))

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--a"); ap.add_argument("--b"); ap.add_argument("--rounds", type=int, default=2)
    ap.add_argument("--libs", help="comma-separated list of more than two builds (replaces --a/--b); the first is the baseline")
    ap.add_argument("--reps", type=int, default=3); ap.add_argument("--child", action="store_true")
    a = ap.parse_args()
    if a.child:
        child(a.reps); sys.exit(0)
    if a.libs:
        libs = a.libs.split(",")
        res = {l: [] for l in libs}
        for r in range(a.rounds):
            for l in libs:
                env = dict(os.environ, SWGPU_LIB=os.path.abspath(l))
                o = subprocess.run([sys.executable, __file__, "--child", "--reps", str(a.reps)], env=env, capture_output=True, text=True)
                if o.returncode:
                    print(o.stderr[-2000:]); sys.exit(1)
                d = json.loads(o.stdout.strip().splitlines()[-1]); res[l].append(d)
                print(json.dumps({"lib": os.path.basename(l), "many_ms": d["many_ms"], "crc": d.get("crc")}), flush=True)
        best = lambda l, lq: min(d["ms"][str(lq)] for d in res[l])
        for lq in LQS:
            base = best(libs[0], lq)
            print(json.dumps({"lq": lq, "base_ms": base} | {os.path.basename(l)[9:-3]: round(best(l, lq) / base, 4) for l in libs[1:]}))
        print(json.dumps({os.path.basename(l): min(d["many_ms"] for d in res[l]) for l in libs}))
        sys.exit(0)
    res = {"a": [], "b": []}
    for r in range(a.rounds):
        for key in ("a", "b"):
            env = dict(os.environ, SWGPU_LIB=os.path.abspath(getattr(a, key)))
            o = subprocess.run([sys.executable, __file__, "--child", "--reps", str(a.reps)], env=env, capture_output=True, text=True)
            if o.returncode:
                print(o.stderr[-2000:]); sys.exit(1)
            d = json.loads(o.stdout.strip().splitlines()[-1]); res[key].append(d)
            print(json.dumps({"arm": key, "many_ms": d["many_ms"], "many_gcups": d["many_gcups"]}), flush=True)
    best = lambda k, lq: min(d["ms"][str(lq)] for d in res[k])
    for lq in LQS:
        print(json.dumps({"lq": lq, "a_ms": best("a", lq), "b_ms": best("b", lq), "b_over_a": round(best("b", lq) / best("a", lq), 4)}))
    print(json.dumps({"many_a": min(d["many_ms"] for d in res["a"]), "many_b": min(d["many_ms"] for d in res["b"])}))
