"""GPU suite: config 5 of BASELINE.json (SURVEY.md 8(d) "C5") at 1/1000 scale -- UniRef50-shaped lengths,
BLOSUM45, 64 batched queries through swgpu_search_many(), the database dealt into shards and the per-shard
top-100 merged with the rule of SURVEY.md A.7.  Parity as the survey specifies for C5: a deterministic sample
(every 64th target) plus every reported hit goes through the scalar int32 oracle."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

K = 100


def test_c5_scaled_sharded_batched(oracle, matrices):
    import swgpu
    from swgpu import synth
    from swgpu.shard import deal_shards, take_shard, merge_topk
    tab = matrices["BLOSUM45"]
    n = 33_000
    codes, off = synth.make_db(46, n, median=230.0, sigma=0.75, lo=11, hi=35213, total=10_000_000, force_max=12_000,
                               sort=False)
    lens = synth.C2_QUERY_LENGTHS
    qs = [synth.make_query(4600 + i, lens[i % len(lens)]) for i in range(64)]
    g = synth.rng(47)
    for i in range(0, n, 97):                                   # related segments so that top hits are real alignments
        q = qs[i % 64]; L = int(off[i + 1] - off[i])
        m = int(min(L - 2, len(q) - 2, g.integers(8, 120)))
        if m < 8:
            continue
        a = int(g.integers(0, len(q) - m)); p = int(g.integers(0, L - m))
        codes[off[i] + p: off[i] + p + m] = q[a:a + m]

    # one GPU holding everything: the reference result for "independent of the number of shards"
    db = swgpu.Database(codes, off)
    scores, hits1, per, bs = db.search_many(qs, tab, -10, -1, k=K)
    db.close()
    assert bs["cells"] == sum(len(q) for q in qs) * int(off[-1])

    # oracle: every 64th target for every 8th query, and all reported hits of three queries
    sample = np.arange(0, n, 64)
    for qi in range(0, 64, 8):
        for t in sample:
            want = oracle.score32(qs[qi], codes[off[t]:off[t + 1]], tab, -10, -1)
            assert scores[qi][t] == want, (qi, t)
    for qi in (0, 19, 63):
        for h in hits1[qi][:25]:
            e = h["entry"]
            s, c = oracle.end_start(qs[qi], codes[off[e]:off[e + 1]], tab, -10, -1)
            assert (h["score"], h["q_start"], h["q_end"], h["t_start"], h["t_end"]) == (s, *c)

    # the same search over 3 shards (length-balanced deal), per-shard top-k merged: identical hit lists
    shards = deal_shards(off, 3)
    res = [(sum(int(off[e + 1] - off[e]) for e in s)) for s in shards]
    assert max(res) - min(res) <= 35213                          # residue balance within one target
    per_shard = []
    for entries in shards:
        c, o = take_shard(codes, off, entries)
        d = swgpu.Database(c, o)
        _, hs, _, _ = d.search_many(qs, tab, -10, -1, k=K, want_scores=False)
        d.close()
        per_shard.append([[(h["score"], int(entries[h["entry"]])) for h in hq] for hq in hs])
    for qi in range(64):
        merged = merge_topk([np.array(ps[qi], dtype=np.int64) for ps in per_shard], K)
        want = [(h["score"], h["entry"]) for h in hits1[qi]]
        assert [tuple(int(x) for x in r) for r in merged] == want, qi
