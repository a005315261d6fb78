"""GPU suite (-m gpu): the BENCHMARK configuration and the other BASELINE configs at their named sizes, replayed against
golden vectors frozen from the unmodified reference (oracle/_ref) and the int32 scalar oracle in the dev container
(tests/golden/make_golden_r2.py).  The databases are regenerated here from their seeds; everything goes through the
C-ABI (swgpu_search_many / swgpu_search)."""
import numpy as np
import pytest

from conftest import golden

pytestmark = pytest.mark.gpu
C2_LQ = (144, 375, 1000, 2005, 5478)


@pytest.fixture(scope="module")
def swgpu_mod():
    import swgpu
    assert swgpu.lib().swgpu_device_count() >= 1, "GPU suite needs a CUDA device; there is no CPU fallback"
    return swgpu


def test_config2_benchmark_database_full_size(swgpu_mod, oracle, matrices):
    """C2 exactly as bench.py builds it (seed 43, 570 000 targets / ~205 M residues, default packing: 2 304-column streams,
    taper, the 35 213-column stream in its own launch), five query lengths through the pipelined many-query entry point
    with top-100 + coordinates: every 512-th batch of 16 targets against the reference's own kernel output, and every
    reported hit against the int32 oracle (the reference's "compare every entry" loop, test/tests.py:105-130, on a sample)."""
    from swgpu import synth
    g = golden("c2_sample.npz")
    tab = matrices["BLOSUM62"]
    codes, off = synth.config2()
    assert len(off) - 1 == int(g["n_targets"]) and int(off[-1]) == int(g["n_residues"])        # same database as the golden
    queries = [synth.make_query(100 + synth.C2_QUERY_LENGTHS.index(lq), lq) for lq in C2_LQ]
    db = swgpu_mod.Database(codes, off)
    scores, hits, per, bs = db.search_many(queries, tab, -10, -1, k=100)
    idx = g["idx"]
    for qi, lq in enumerate(C2_LQ):
        want = g[f"ref16_{lq}"].astype(np.int32)
        got = scores[qi][idx]
        bad = np.nonzero(got != want)[0]
        assert bad.size == 0, (lq, idx[bad[:8]], got[bad[:8]], want[bad[:8]])
        assert per[qi]["cells"] == lq * int(off[-1])
        # top-100: sorted (score desc, entry asc), consistent with the score vector, and each hit == int32 oracle
        hl = hits[qi]
        assert len(hl) == 100
        order = np.lexsort((np.arange(len(off) - 1), -scores[qi].astype(np.int64)))[:100]
        assert [h["entry"] for h in hl] == order.tolist()
        for h in hl[:: 10 if lq > 1000 else 4]:
            e = h["entry"]
            s, c = oracle.end_start(queries[qi], codes[off[e]:off[e + 1]], tab, -10, -1)
            assert (h["score"], h["q_start"], h["q_end"], h["t_start"], h["t_end"]) == (s, *c), (lq, h)
    # the single-query path gives the same vector as the pipelined one
    got1, st = db.search(queries[3], tab, -10, -1)
    assert np.array_equal(got1, scores[3]) and st["passes"] > 1
    db.close()


def test_config3_overflow_stress_full_size(swgpu_mod, matrices):
    """C3 at SURVEY.md 8(d)'s size: 4 096 targets, 6 000-aa low-complexity query, PAM250.  GPU == int32 oracle for all
    4 096; GPU == the reference wherever the true score fits its int16; the number of 32-bit-rescored targets is at least
    the number whose score exceeds the 16-bit path's exact range."""
    from swgpu import synth
    g = golden("c3_full.npz")
    q, codes, off = synth.config3()
    assert int(off[-1]) == int(g["n_residues"])
    true32, ref16 = g["true32"], g["ref16"].astype(np.int32)
    db = swgpu_mod.Database(codes, off)
    got, st = db.search(q, matrices["PAM250"], -10, -1)
    bad = np.nonzero(got != true32)[0]
    assert bad.size == 0, (bad[:8], got[bad[:8]], true32[bad[:8]])
    fits = true32 <= 32767
    assert int((~fits).sum()) >= 2000 and np.array_equal(got[fits], ref16[fits])
    n_beyond = int((true32 > 65535 - 129 - (128 + 11 + 1)).sum())       # beyond the biased-u16 range of K1S
    assert n_beyond >= 400 and st["n_rescored"] >= n_beyond, (st["n_rescored"], n_beyond)
    db.close()
    # the same through the 32-bit kernels alone
    db32 = swgpu_mod.Database(codes, off, long_threshold=1)
    got32, st32 = db32.search(q, matrices["PAM250"], -10, -1)
    assert np.array_equal(got32, true32) and st32["n_rescored"] == len(off) - 1
    db32.close()


def test_config4_oracle_sample_of_121(swgpu_mod, matrices):
    """C4 at full size against the int32 oracle on 121 targets: the first 64 (all 16 titin-length ones, the 90 %-identity
    copy, the 48 log-uniform 8-35 k ones) and every 17-th of the rest."""
    from swgpu import synth
    g = golden("c4_sample.npz")
    q, codes, off = synth.config4()
    assert int(off[-1]) == int(g["n_residues"])
    db = swgpu_mod.Database(codes, off)
    got, st = db.search(q, matrices["BLOSUM62"], -10, -1)
    db.close()
    idx, want = g["idx"], g["true32"]
    assert len(idx) >= 64 and np.array_equal(idx[:64], np.arange(64))
    bad = np.nonzero(got[idx] != want)[0]
    assert bad.size == 0, (idx[bad[:8]], got[idx][bad[:8]], want[bad[:8]])
    assert int(want.max()) > 65535 and st["n_rescored"] >= 1
