"""GPU suite: swgpu_search_many() -- several queries in flight on separate slots -- must return exactly what
one swgpu_search() + swgpu_topk() per query returns (and what the oracle says)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _planted_db(seed, n, qs):
    from swgpu import synth
    codes, off = synth.make_db(seed, n, median=180.0, lo=1, hi=5000, force_max=5000)
    g = synth.rng(seed + 1)
    for t in range(0, n, 5):
        L = int(off[t + 1] - off[t])
        q = qs[t % len(qs)]
        if L < 40 or len(q) < 40:
            continue
        m = int(min(L - 5, len(q) - 5, g.integers(10, 300)))
        a = int(g.integers(0, len(q) - m)); p = int(g.integers(0, L - m))
        codes[off[t] + p: off[t] + p + m] = q[a:a + m]
    return codes, off


def test_many_equals_single_and_oracle(oracle, matrices):
    import swgpu
    from swgpu import synth
    tab = matrices["BLOSUM62"]
    lens = [144, 33, 375, 1, 1000, 257, 2005, 64, 567]          # more queries than slots, all K1S plans incl. multi-stripe
    qs = [synth.make_query(500 + i, L) for i, L in enumerate(lens)]
    codes, off = _planted_db(71, 4000, qs)
    db = swgpu.Database(codes, off, entry_base=7)
    k = 50
    scores, hits, per, bs = db.search_many(qs, tab, -10, -1, k=k)
    assert bs["slots"] == 3 and bs["cells"] == sum(lens) * int(off[-1]) and bs["ms_total"] > 0
    for i, q in enumerate(qs):
        want = oracle.search32(q, codes, off, tab, -10, -1)
        assert np.array_equal(scores[i], want), f"query {i} (Lq={lens[i]})"
        s1, _ = db.search(q, tab, -10, -1)
        assert np.array_equal(s1, want)
        h1 = db.topk(k)
        assert hits[i] == h1, f"query {i}: hits differ between search_many and search+topk"
        assert per[i]["cells"] == lens[i] * int(off[-1])
    # coordinates of a few hits against the oracle
    for i in (0, 4, 6):
        for h in hits[i][:3]:
            e = h["entry"] - 7
            s, c = oracle.end_start(qs[i], codes[off[e]:off[e + 1]], tab, -10, -1)
            assert (h["score"], h["q_start"], h["q_end"], h["t_start"], h["t_end"]) == (s, *c)
    db.close()


def test_many_pinned_outputs_rescoring_and_no_topk(oracle, matrices):
    import torch
    import swgpu
    from swgpu import synth
    tab = matrices["PAM250"]
    g = synth.rng(91)
    # low-complexity query whose copies saturate the 16-bit pass (C3-style), next to ordinary queries
    qw = g.choice(np.array([23, 3, 25, 6], dtype=np.int8), size=6000, p=[0.4, 0.2, 0.2, 0.2]).astype(np.int8)
    qs = [qw, synth.make_query(92, 300), qw[:5200].copy(), synth.make_query(93, 90)]
    codes, off = synth.make_db(94, 600, median=200.0, lo=5, hi=3000)
    extra = [qw[100:5600], qw[:5900], qw[2000:4500]]
    codes = np.concatenate([codes] + extra)
    off = np.concatenate([off, off[-1] + np.cumsum([len(x) for x in extra])])
    db = swgpu.Database(codes, off)
    n = len(off) - 1
    out = [torch.empty(n, dtype=torch.int32).pin_memory().numpy() for _ in qs]
    scores, hits, per, bs = db.search_many(qs, tab, -10, -1, k=0, out=out)
    assert hits is None
    assert per[0]["n_rescored"] >= 2 and per[1]["n_rescored"] == 0
    assert bs["n_rescored"] == sum(p["n_rescored"] for p in per)
    for i, q in enumerate(qs):
        assert np.array_equal(scores[i], oracle.search32(q, codes, off, tab, -10, -1)), i
    db.close()


def test_many_degenerate(matrices):
    import swgpu
    from swgpu import synth
    tab = matrices["BLOSUM62"]
    codes, off = synth.make_db(5, 40, median=50.0, lo=1, hi=200)
    db = swgpu.Database(codes, off)
    scores, hits, per, bs = db.search_many([], tab, -10, -1, k=3)
    assert scores == [] and hits == [] and bs["cells"] == 0
    # empty query among real ones; k larger than the database
    qs = [synth.make_query(1, 50), np.zeros(0, dtype=np.int8), synth.make_query(2, 20)]
    scores, hits, per, bs = db.search_many(qs, tab, -10, -1, k=100)
    assert (scores[1] == 0).all() and len(hits[1]) == 40 and all(h["q_start"] == -1 for h in hits[1])
    assert [h["entry"] for h in hits[1]] == list(range(40))
    assert len(hits[0]) == 40 and hits[0][0]["score"] == int(scores[0].max())
    with pytest.raises(swgpu.SwgpuError):
        db.search_many([np.array([40], dtype=np.int8)], tab, -10, -1)
    # the handle still works after an argument error
    s2, _ = db.search(qs[0], tab, -10, -1)
    assert np.array_equal(s2, scores[0])
    db.close()
