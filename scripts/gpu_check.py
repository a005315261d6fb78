"""Quick GPU parity + timing check against the oracle (dev tool; tests/ holds the real suite)."""
import sys, time, json
sys.path.insert(0, '/root/repo/seq-align-gpu_b200'); sys.path.insert(0, '/root/repo')
import numpy as np
import swgpu
from swgpu import synth
from oracle import oracle as O

tab, diag = synth.load_matrix('/root/repo/scoring/BLOSUM62.txt')
q, codes, off = synth.config1()
n = len(off) - 1
t0 = time.time(); db = swgpu.Database(codes, off); print("pack+upload s", time.time() - t0, "bytes", db.device_bytes)
for lq in (360, 144, 33, 1):
    qq = q[:lq]
    s, st = db.search(qq, tab, -10, -1)
    ref = O.search32(qq, codes, off, tab, -10, -1)
    bad = np.nonzero(s != ref)[0]
    print("C1 lq", lq, "mismatches", bad.size, "of", n, "| GCUPS", st['cells'] / st['ms_total'] / 1e6, json.dumps(st))
    if bad.size:
        print("  first bad", bad[:10], s[bad[:10]], ref[bad[:10]], (off[1:] - off[:-1])[bad[:10]])
# force everything through K3 (long_threshold tiny) on a subset
sub = 600
dbl = swgpu.Database(codes[:off[sub]], off[:sub + 1], long_threshold=1)
s, st = dbl.search(q, tab, -10, -1)
ref = O.search32(q, codes[:off[sub]], off[:sub + 1], tab, -10, -1)
print("K3-only mismatches", int((s != ref).sum()), "GCUPS", st['cells'] / st['ms_total'] / 1e6, st['n_long'])
# repeat timing
for i in range(3):
    st = db.search_resident(q, tab, -10, -1)
    print("resident run", i, "ms", st['ms_total'], "GCUPS", st['cells'] / st['ms_total'] / 1e6)
