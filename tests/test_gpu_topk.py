"""GPU suite: device-side top-k and linear-space coordinates against the oracle's rule (SURVEY.md A.7;
PARITY UNPINNED by the reference, which keeps only a per-target maximum)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def expect_topk(scores, k, base=0):
    order = np.lexsort((np.arange(scores.size), -scores.astype(np.int64)))[:k]
    return [(int(scores[i]), int(i) + base) for i in order]


@pytest.mark.parametrize("k", [1, 5, 100, 1000, 5000])
def test_topk_order_and_coordinates(oracle, matrices, k):
    import swgpu
    from swgpu import synth
    tab = matrices["BLOSUM62"]
    codes, off = synth.make_db(61, 3000, median=150.0, lo=1, hi=2500)
    q = synth.make_query(62, 411)
    g = synth.rng(63)
    for t in range(0, 3000, 7):                       # plant related segments so hits are real alignments
        L = int(off[t + 1] - off[t])
        if L < 30:
            continue
        a = int(g.integers(0, 380)); n = int(min(L - 5, g.integers(10, 200), 411 - a))
        p = int(g.integers(0, L - n))
        codes[off[t] + p: off[t] + p + n] = q[a:a + n]
    db = swgpu.Database(codes, off, entry_base=1000)
    scores, _ = db.search(q, tab, -10, -1)
    assert np.array_equal(scores, oracle.search32(q, codes, off, tab, -10, -1))
    hits = db.topk(k)
    want = expect_topk(scores, min(k, 3000), base=1000)
    assert [(h["score"], h["entry"]) for h in hits] == want       # many ties at low scores: entry asc breaks them
    step = max(1, len(hits) // 60)
    for h in hits[::step]:
        i = h["entry"] - 1000
        s, c = oracle.end_start(q, codes[off[i]:off[i + 1]], tab, -10, -1)
        assert (h["score"], h["q_start"], h["q_end"], h["t_start"], h["t_end"]) == (s, *c), h
    db.close()


def test_topk_long_query_multi_stripe_and_zero_scores(oracle, matrices):
    import swgpu
    from swgpu import synth
    tab = matrices["PAM250"]
    g = synth.rng(70)
    q = synth.make_query(71, 1500)                       # > 256 rows: several stripes in the coordinate kernel
    parts = [synth.residues(g, int(L)) for L in g.integers(1, 4000, size=40)]
    parts[3] = np.concatenate([parts[3][:100], q[200:1400], parts[3][100:]])
    parts[9] = q[::-1].copy()
    parts += [np.zeros(0, np.int8)] * 3                    # empty targets score 0 and carry no coordinates
    off = np.zeros(len(parts) + 1, np.int64); np.cumsum([p.size for p in parts], out=off[1:])
    codes = np.concatenate(parts)
    db = swgpu.Database(codes, off, long_threshold=1000)
    scores, _ = db.search(q, tab, -12, -2)
    hits = db.topk(len(parts))
    assert [(h["score"], h["entry"]) for h in hits] == expect_topk(scores, len(parts))
    for h in hits:
        i = h["entry"]
        s, c = oracle.end_start(q, codes[off[i]:off[i + 1]], tab, -12, -2)
        assert (h["score"], h["q_start"], h["q_end"], h["t_start"], h["t_end"]) == (s, *c), h
    assert hits[-1]["score"] == 0 and hits[-1]["q_end"] == -1


def test_topk_coordinates_with_ties_and_early_stop(oracle, matrices):
    """The coordinate passes stop at the first column that holds the hit's (known) score.  Ties are the dangerous case: the
    same maximal score at several columns and in several stripes of query rows -- the FIRST column in the reference's scan
    order must win, whichever stripe gets there first.  Targets hold two or three exact copies of query
    segments (equal maxima far apart, in different 256-row stripes), a copy at the very first and at the very last column,
    and long background; the same hits come out with the early stop switched off (SWGPU_TOPK_NO_EARLY_STOP)."""
    import os
    import swgpu
    from swgpu import synth
    tab = matrices["BLOSUM62"]
    g = synth.rng(170)
    q = synth.make_query(171, 900)                           # 4 stripes of 256 rows in the coordinate kernel
    bgr = lambda n: synth.residues(g, int(n))
    parts = [
        np.concatenate([bgr(700), q[100:300], bgr(1500), q[100:300], bgr(900)]),               # the same segment twice: equal maxima
        np.concatenate([bgr(300), q[600:850], bgr(50), q[20:270], bgr(2000), q[600:850]]),     # equal-length copies from different stripes
        np.concatenate([q[400:700], bgr(5000)]),                                               # hit at the first columns of a long target
        np.concatenate([bgr(6000), q[0:250]]),                                                 # hit at the last columns
        np.concatenate([q[300:420]] * 5),                                                      # five abutting copies
        bgr(9000), bgr(40), q.copy(), q[::-1].copy(),
    ] + [bgr(L) for L in g.integers(1, 1200, size=40)]
    off = np.zeros(len(parts) + 1, np.int64); np.cumsum([p.size for p in parts], out=off[1:])
    codes = np.concatenate(parts)
    db = swgpu.Database(codes, off)
    results = []
    for env in (None, "1"):
        if env: os.environ["SWGPU_TOPK_NO_EARLY_STOP"] = env
        else: os.environ.pop("SWGPU_TOPK_NO_EARLY_STOP", None)
        for lq in (900, 130):
            db.search(q[:lq], tab, -10, -1)
            hits = db.topk(len(parts))
            results.append((env, lq, hits, db.last_topk_ms))
            for h in hits:
                i = h["entry"]
                s, c = oracle.end_start(q[:lq], codes[off[i]:off[i + 1]], tab, -10, -1)
                assert (h["score"], h["q_start"], h["q_end"], h["t_start"], h["t_end"]) == (s, *c), (env, lq, h, c)
    os.environ.pop("SWGPU_TOPK_NO_EARLY_STOP", None)
    assert results[0][2] == results[2][2] and results[1][2] == results[3][2]
    print(f"top-{len(parts)} coordinates, Lq=900: {results[0][3]:.2f} ms with the early stop, {results[2][3]:.2f} ms without")
    db.close()


def test_topk_row_bound_of_the_forward_pass_with_ties_across_stripes(oracle, matrices):
    """The forward coordinate pass stops at the last query row that can hold the hit's score (k1s_finalize records the last
    lane slot holding the maximum; X pairs keep one set of slots per stripe).  The query carries the same 200-residue segment
    twice, 1 100 rows apart -- in different stripes of the two-stripe 16-bit plan -- so a target holding it once has two equal
    maxima in different row bands: the bound must keep both, the first one in scan order must win, and the hits must be the
    same with the bound switched off (SWGPU_TOPK_NO_ROW_BOUND).  The database is small: every stream is an X item."""
    import os
    import swgpu
    from swgpu import synth
    tab = matrices["BLOSUM62"]
    g = synth.rng(270)
    q0 = synth.make_query(271, 900)
    bgr = lambda n: synth.residues(g, int(n))
    q = np.concatenate([q0, bgr(300), q0[100:300], bgr(700)])           # 2 100 rows: q0[100:300] at rows 100.. and 1 200..
    parts = [
        np.concatenate([bgr(500), q0[100:300], bgr(800)]),              # one copy: equal maxima in both stripes' bands
        np.concatenate([bgr(200), q0[100:300], bgr(900), q0[100:300], bgr(100)]),
        np.concatenate([bgr(1200), q[1900:2090], bgr(30)]),             # a hit in the last rows
        np.concatenate([q0[0:150], bgr(2000)]),                         # a hit in the first rows
        bgr(3000), q0[100:300].copy(),
    ] + [bgr(L) for L in g.integers(1, 900, size=30)]
    off = np.zeros(len(parts) + 1, np.int64); np.cumsum([p.size for p in parts], out=off[1:])
    codes = np.concatenate(parts)
    db = swgpu.Database(codes, off)
    out = []
    for env in (None, "1"):
        if env: os.environ["SWGPU_TOPK_NO_ROW_BOUND"] = env
        else: os.environ.pop("SWGPU_TOPK_NO_ROW_BOUND", None)
        _, st = db.search(q, tab, -10, -1)
        assert st["passes"] >= 2
        hits = db.topk(len(parts))
        out.append(hits)
        for h in hits:
            i = h["entry"]
            s, c = oracle.end_start(q, codes[off[i]:off[i + 1]], tab, -10, -1)
            assert (h["score"], h["q_start"], h["q_end"], h["t_start"], h["t_end"]) == (s, *c), (env, h, c)
    os.environ.pop("SWGPU_TOPK_NO_ROW_BOUND", None)
    assert out[0] == out[1]
    first = [h for h in out[0] if h["entry"] == 0][0]
    assert first["q_end"] < 400                                          # of the two equal maxima the one in the upper rows wins
    db.close()
