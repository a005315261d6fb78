"""The smallest case that reaches every kernel family, for compute-sanitizer (scripts/gpu_sanitize.sh <tool>):
K1S single-stripe (8 and 32 lanes per pair, the long-stream launch), K1S multi-stripe with X items (release/acquire progress
words between CTAs), k1s_finalize, K3 / K3X (32-bit; K3X's tag-in-data hand-over between warps), top-k selection +
coordinate passes, tb_fill_trace (traceback), k3_three (positive gap scores), swgpu_fill_batch16, the device generator.
Every result is checked against the oracle, so a sanitizer run is also a parity run."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "seq-align-gpu_b200"))
import swgpu
from swgpu import synth
from oracle import oracle as O

tab, _ = synth.load_matrix(os.path.join(ROOT, "scoring", "BLOSUM62.txt"))
g = synth.rng(99)
lens = np.concatenate([[4500, 2100, 0], synth.lognormal_lengths(g, 90, lo=1, hi=700)])
off = np.zeros(len(lens) + 1, dtype=np.int64); np.cumsum(lens, out=off[1:])
codes = synth.residues(g, off[-1])
q = synth.make_query(98, 1100)
codes[off[1] + 100: off[1] + 1000] = q[50:950]
checks = 0
for kw, cases in (({}, [(150, -10, -1), (700, -10, -1), (1100, -10, -1), (1100, -7, -2), (90, 1, 1)]),
                  ({"long_threshold": 1}, [(300, -10, -1)])):
    db = swgpu.Database(codes, off, **kw)
    for lq, go, ge in cases:
        got, st = db.search(q[:lq], tab, go, ge)
        want = O.search32(q[:lq], codes, off, tab, go, ge)
        assert np.array_equal(got, want), (kw, lq, go, ge, int((got != want).sum()))
        checks += 1
        if go < 0:
            hits = db.topk(4)
            for h, cg in zip(hits, db.traceback(q[:lq], tab, go, ge, hits)):
                e = int(h["entry"]); t = codes[off[e]:off[e + 1]]
                s, c = O.end_start(q[:lq], t, tab, go, ge)
                assert (h["score"], h["q_start"], h["q_end"], h["t_start"], h["t_end"]) == (s, *c), h
                if s > 0:
                    assert O.cigar_score(cg, q[:lq][c[0]:c[1] + 1], t[c[2]:c[3] + 1], tab, go, ge) == s
    db.close()
batch = np.full((60, 16), 31, dtype=np.int8)
for l in range(16):
    batch[: 60 - 3 * l, l] = synth.residues(g, 60 - 3 * l)
got = swgpu.fill_batch16(q[:70], batch.reshape(-1), 60, 16, tab, -10, -1)
assert list(got) == [O.score16_wrap(q[:70], batch[: 60 - 3 * l, l], 60, tab, -10, -1) for l in range(16)]
sl = synth.c5_lengths(0.00001)[:60]
sdb = swgpu.Database.synthetic(sl, synth.C5_SEED)
assert np.array_equal(sdb.fetch_target(7), synth.synth_target(synth.C5_SEED, 7, int(sl[7])))
sdb.close()
print(f"sanitize_case: {checks} searches + top-k + traceback + batch16 + generator OK")
