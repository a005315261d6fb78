"""Two k3x launches for one ncu capture (scripts/gpu_r2_ncu_k3x.sh): ONE 35 213 x 35 213 pair (138 stripes on an otherwise idle
machine: the latency case), then 64 such targets (8 832 items on 2 368 warps: the throughput case).  long_threshold = 1 sends
every target through the 32-bit path."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "seq-align-gpu_b200"))
import numpy as np
import swgpu
from swgpu import synth
tab, _ = synth.load_matrix(os.path.join(ROOT, "scoring", "BLOSUM62.txt"))
lt = 35213
q = synth.make_query(7, lt)
for nt in (1, 64):
    g = synth.rng(5)
    off = np.arange(nt + 1, dtype=np.int64) * lt
    db = swgpu.Database(synth.residues(g, off[-1]), off, long_threshold=1)
    st = db.search_resident(q, tab, -10, -1)
    print(json.dumps({"targets": nt, "ms_rescore": round(st["ms_rescore"], 3), "gcups32": round(lt * nt * lt / st["ms_rescore"] / 1e6, 1)}), flush=True)
    db.close()
