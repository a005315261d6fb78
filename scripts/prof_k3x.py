"""K3X in isolation: a few titin-length targets forced through the 32-bit wavefront kernel, several query lengths."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "seq-align-gpu_b200"))
import numpy as np
import swgpu
from swgpu import synth
tab, _ = synth.load_matrix(os.path.join(ROOT, "scoring", "BLOSUM62.txt"))
g = synth.rng(1)
nt = int(sys.argv[1]) if len(sys.argv) > 1 else 2
lt = int(sys.argv[2]) if len(sys.argv) > 2 else 35213
parts = [synth.residues(g, lt) for _ in range(nt)]
off = np.zeros(nt + 1, np.int64); np.cumsum([p.size for p in parts], out=off[1:])
db = swgpu.Database(np.concatenate(parts), off, long_threshold=1)
for lq in (128, 256, 1024, 2048, 4096, 8192, 35213):
    q = synth.make_query(5, lq)
    db.search_resident(q, tab, -10, -1)
    st = db.search_resident(q, tab, -10, -1)
    stripes = -(-lq // 128); iters = -(-stripes // 8)
    steps = iters * (lt + 31)
    print(json.dumps({"lq": lq, "ms_rescore": round(st["ms_rescore"], 3), "ms_inter": round(st["ms_inter"], 3), "super_iters": iters,
                      "ns_per_step": round(st["ms_rescore"] * 1e6 / steps, 1), "cycles_per_step": round(st["ms_rescore"] * 1e6 / steps * 1.965, 1),
                      "gcups32": round(lq * nt * lt / st["ms_rescore"] / 1e6, 2)}))
