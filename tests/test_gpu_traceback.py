"""GPU suite: alignment paths (CIGAR) of reported hits, swgpu_traceback().  PARITY UNPINNED by the reference (it keeps
only a score per target); pinned against this repo's scalar oracle (oracle/sw_oracle.c::swo_traceback) operation for
operation, and checked independently: a path must consume exactly the hit's sub-sequences, start and end with an
aligned pair, and re-score to the hit's score."""
import re

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def planted(seed, n, q, indel_every=9):
    """database with noisy, gapped copies of pieces of q planted in every third target"""
    from swgpu import synth
    codes, off = synth.make_db(seed, n, median=400.0, lo=20, hi=3000)
    g = synth.rng(seed + 1)
    for t in range(0, n, 3):
        L = int(off[t + 1] - off[t])
        m = int(min(L - 10, len(q) - 1, g.integers(30, 900)))
        if m < 30:
            continue
        a = int(g.integers(0, len(q) - m))
        piece = list(q[a:a + m])
        out = []
        for k, c in enumerate(piece):                     # substitutions, deletions, insertions
            u = g.random()
            if u < 0.04:
                out.append(int(synth.residues(g, 1)[0]))
            elif u < 0.04 + 1.0 / (indel_every * 4):
                continue
            elif u < 0.04 + 2.0 / (indel_every * 4):
                out.extend([int(c)] + [int(x) for x in synth.residues(g, int(g.integers(1, 6)))])
            else:
                out.append(int(c))
        out = np.array(out[: L - 2], dtype=np.int8)
        p = int(g.integers(0, L - len(out)))
        codes[off[t] + p: off[t] + p + len(out)] = out
    return codes, off


@pytest.mark.parametrize("lq,matrix,gaps", [(300, "BLOSUM62", (-10, -1)), (1100, "BLOSUM45", (-12, -2)), (700, "PAM250", (-5, -3))])
def test_traceback_equals_oracle_and_rescans(oracle, matrices, lq, matrix, gaps):
    import swgpu
    from swgpu import synth
    tab = matrices[matrix]
    q = synth.make_query(900 + lq, lq)
    codes, off = planted(lq, 600, q)
    db = swgpu.Database(codes, off, entry_base=11)
    db.search_resident(q, tab, *gaps)
    hits = db.topk(80)
    cigars = db.traceback(q, tab, gaps[0], gaps[1], hits)
    assert len(cigars) == len(hits)
    multi_stripe = 0
    for h, cg in zip(hits, cigars):
        e = h["entry"] - 11
        t = codes[off[e]:off[e + 1]]
        qs, ts = q[h["q_start"]:h["q_end"] + 1], t[h["t_start"]:h["t_end"] + 1]
        want, sc = oracle.traceback(qs, ts, tab, *gaps)
        assert sc == h["score"]
        assert cg == want, (h, cg, want)
        assert oracle.cigar_score(cg, qs, ts, tab, *gaps) == h["score"]
        ops = re.findall(r"(\d+)([MID])", cg)
        assert ops[0][1] == "M" and ops[-1][1] == "M"
        multi_stripe += len(qs) > 256
    assert any("I" in c for c in cigars) and any("D" in c for c in cigars)
    if lq > 600:
        assert multi_stripe > 0                            # rectangles taller than one 256-row stripe
    db.close()


def test_traceback_after_search_many_zero_scores_and_errors(oracle, matrices):
    import swgpu
    from swgpu import synth
    tab = matrices["BLOSUM62"]
    qs = [synth.make_query(31, 200), synth.make_query(32, 450)]
    codes, off = planted(77, 200, qs[1])
    db = swgpu.Database(codes, off)
    _, hits, _, _ = db.search_many(qs, tab, -10, -1, k=10, want_scores=False)
    for q, hq in zip(qs, hits):
        cg = db.traceback(q, tab, -10, -1, hq)
        for h, c in zip(hq, cg):
            t = codes[off[h["entry"]]:off[h["entry"] + 1]]
            assert oracle.cigar_score(c, q[h["q_start"]:h["q_end"] + 1], t[h["t_start"]:h["t_end"] + 1], tab, -10, -1) == h["score"]
    # a hit without coordinates (score 0) gets an empty path; no hits at all is fine
    zero = dict(score=0, q_start=-1, q_end=-1, t_start=-1, t_end=-1, entry=3)
    assert db.traceback(qs[0], tab, -10, -1, [zero, hits[0][0]])[0] == ""
    assert db.traceback(qs[0], tab, -10, -1, []) == []
    # the wrong scoring parameters are detected (the rectangle no longer scores hit.score)
    with pytest.raises(swgpu.SwgpuError):
        db.traceback(qs[1], tab, -3, -1, hits[1][:3])
    bad = dict(hits[1][0]); bad["entry"] = 10**6
    with pytest.raises(swgpu.SwgpuError):
        db.traceback(qs[1], tab, -10, -1, [bad])
    db.close()
