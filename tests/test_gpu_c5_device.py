"""GPU suite: config 5 of BASELINE.json (SURVEY.md 8(d) "C5") with the database GENERATED ON THE DEVICE
(swgpu_db_create_synthetic): UniRef50-shaped lengths, BLOSUM45, 64 batched queries, k = 100, shards dealt by
length and merged with the rule of SURVEY.md A.7.

The default size is 1/100 of C5 (100 M residues, 328 k targets); SWGPU_C5_SCALE=1.0 runs the named size (10 G
residues, 32.8 M targets, on one B200) -- that run is recorded in profiles/ (r1f_c5_full.txt).  The checks do not
depend on the size:
  * generator: residues fetched back from the packed database == the host mirror (swgpu.synth.synth_target);
  * parity as the survey specifies for C5: a deterministic sample (every STRIDE-th target) and the reported hits go
    through the scalar int32 oracle on host-regenerated targets;
  * every query carries a 90 %-identity copy of one database target: that target must be its best hit;
  * top-k lists are sorted (score desc, entry asc) and consistent with the full score vector;
  * the same search over dealt shards (global entry ids) + merge gives the identical hit lists.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

K = 100
SCALE = float(os.environ.get("SWGPU_C5_SCALE", "0.01"))


def test_c5_device_generated(oracle, matrices):
    import swgpu
    from swgpu import synth
    from swgpu.shard import merge_topk
    tab = matrices["BLOSUM45"]
    spec = synth.synth_spec()
    lens = synth.c5_lengths(SCALE)
    n = len(lens)
    qs, planted = synth.c5_queries(lens, 64, spec=spec)
    target = lambda e: synth.synth_target(synth.C5_SEED, e, int(lens[e]), spec)

    db = swgpu.Database.synthetic(lens, synth.C5_SEED, spec=spec)
    assert db.n_residues == int(lens.astype(np.int64).sum())

    # generator == host mirror (first, last, the longest, a spread of others; includes odd/even pair members)
    probe = sorted({0, 1, 2, n - 1, n - 2, int(np.argmax(lens)), int(np.argmin(lens))} | set(range(3, n, max(1, n // 97))))
    for e in probe:
        got = db.fetch_target(e)
        assert got.shape[0] == lens[e] and np.array_equal(got, target(e)), e

    _, hits, per, bs = db.search_many(qs, tab, -10, -1, k=K, want_scores=False)
    assert bs["cells"] == sum(len(q) for q in qs) * int(db.n_residues)
    print(f"\nC5 scale {SCALE}: {n} targets, {db.n_residues} residues, 64 queries: {bs['ms_total']:.1f} ms, "
          f"{bs['cells'] / bs['ms_total'] / 1e6:.0f} GCUPS, {bs['n_rescored']} rescored, {bs['kernel_launches']} launches")

    for qi in range(64):
        hq = hits[qi]
        assert len(hq) == K
        keys = [(-h["score"], h["entry"]) for h in hq]
        assert keys == sorted(keys), qi
        # the planted target is the best hit, with the oracle's score
        assert hq[0]["entry"] == planted[qi], (qi, hq[0], planted[qi])
        assert hq[0]["score"] == oracle.score32(qs[qi], target(planted[qi]), tab, -10, -1), qi

    # reported hits: score and coordinates against the oracle (cost bounded by the shorter queries)
    for qi in (0, 5, 21, 43, 63):
        for h in hits[qi][:12]:
            e = h["entry"]
            s, c = oracle.end_start(qs[qi], target(e), tab, -10, -1)
            assert (h["score"], h["q_start"], h["q_end"], h["t_start"], h["t_end"]) == (s, *c), (qi, h)
    for qi in (1, 30):
        for h in hits[qi]:
            assert h["score"] == oracle.score32(qs[qi], target(h["entry"]), tab, -10, -1), (qi, h)

    # deterministic sample of ALL scores of three queries (short, single-stripe limit, multi-stripe)
    stride = max(64, n // 400)
    sample = np.arange(0, n, stride)
    for qi in (0, 9, 12):
        scores, st = db.search(qs[qi], tab, -10, -1)
        for e in sample:
            assert scores[e] == oracle.score32(qs[qi], target(int(e)), tab, -10, -1), (qi, int(e))
        # the hit list is the head of the full score vector under (score desc, entry asc)
        top = np.lexsort((np.arange(n), -scores.astype(np.int64)))[:K]
        assert [int(t) for t in top] == [h["entry"] for h in hits[qi]], qi
    db.close()

    if SCALE > 0.05:
        return                                               # the sharded re-run doubles the GPU time; sized for the default scale
    # two shards dealt longest-first round-robin, global entry ids: identical merged hit lists
    order = np.argsort(-lens.astype(np.int64), kind="stable")
    per_shard = []
    for r in range(2):
        entries = np.sort(order[r::2])
        d = swgpu.Database.synthetic(lens[entries], synth.C5_SEED, entry_ids=entries, spec=spec)
        assert np.array_equal(d.fetch_target(len(entries) // 2), target(int(entries[len(entries) // 2])))
        _, hs, _, _ = d.search_many(qs, tab, -10, -1, k=K, want_scores=False)
        d.close()
        per_shard.append([[(h["score"], int(entries[h["entry"]])) for h in hq] for hq in hs])
    for qi in range(64):
        merged = merge_topk([np.array(ps[qi], dtype=np.int64) for ps in per_shard], K)
        assert [tuple(int(x) for x in r) for r in merged] == [(h["score"], h["entry"]) for h in hits[qi]], qi


def test_synthetic_rejects_bad_arguments():
    import swgpu
    from swgpu import synth
    cdf, cd, nl = synth.synth_spec()
    bad = cdf.copy(); bad[nl - 1] = 65000
    with pytest.raises(swgpu.SwgpuError):
        swgpu.Database.synthetic(np.array([5, 7], dtype=np.int32), 1, spec=(bad, cd, nl))
    with pytest.raises(swgpu.SwgpuError):
        swgpu.Database.synthetic(np.array([5, -7], dtype=np.int32), 1)
    d = swgpu.Database.synthetic(np.zeros(0, dtype=np.int32), 1)       # empty database is legal
    assert d.n_residues == 0
    d.close()
    d = swgpu.Database.synthetic(np.array([0, 3, 0], dtype=np.int32), 9, entry_base=100)
    assert d.fetch_target(0).size == 0 and np.array_equal(d.fetch_target(1), synth.synth_target(9, 101, 3))
    d.close()
