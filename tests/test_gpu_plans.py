"""GPU suite: EVERY instantiation of the 16-bit systolic kernel K1S and every packing option gives the oracle's scores.

The planner picks one (lanes per pair, rows per lane) plan per query length, so the other suites only reach the
instantiations the planner happens to choose.  Here SWGPU_K1S_PLAN forces each of them in turn:
  * single stripe: lanes 8 / 16 / 32 x rows per lane 8..32 (odd values included), on a database with a few long targets
    (their streams take the separate long-stream launch when lanes < 32);
  * several stripes on a small database (every stream goes through X items: one stripe per work item, global hand-over);
  * several stripes on a database big enough for ordinary items (8 streams x all stripes) next to the X items of its
    long streams, checked on a deterministic sample;
  * the same searches under different packings (stream length, taper on / off): identical scores.
Every instantiation exists three times -- the generic kernel (gap penalties in registers) and the two gap specialisations
(penalties as immediates: --gapopen -10 / -11 with --gapextend -1, csrc/sw_k1s.cuh) -- so every case runs under gap
settings that select each of them; SWGPU_K1S_GENERIC=1 must give the same scores through the generic kernel.
Bit-exact against the scalar int32 oracle (reference recurrence: src/alignment.c:117-176)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MULTI_R = (16, 20, 24, 25, 26, 27, 28, 29, 30, 31, 32)
GAPS = [(-10, -1), (-11, -1), (-7, -2)]          # kernel specialisation 1, specialisation 2, generic


@pytest.fixture()
def plan_env():
    old = os.environ.pop("SWGPU_K1S_PLAN", None)
    os.environ.pop("SWGPU_K1S_GENERIC", None)
    yield
    os.environ.pop("SWGPU_K1S_PLAN", None)
    os.environ.pop("SWGPU_K1S_GENERIC", None)
    if old is not None:
        os.environ["SWGPU_K1S_PLAN"] = old


def small_db(synth, seed=71, n=800):
    codes, off = synth.make_db(seed, n, hi=9000, force_max=9000, sort=False)
    lens = off[1:] - off[:-1]
    return codes, off, lens


@pytest.mark.parametrize("go,ge", GAPS)
def test_every_single_stripe_instantiation(oracle, matrices, plan_env, go, ge):
    import swgpu
    from swgpu import synth
    tab = matrices["BLOSUM62"]
    codes, off, _ = small_db(synth)
    db = swgpu.Database(codes, off)
    want = {}
    for g in (8, 16, 32):
        for R in range(8, 33):
            lq = g * R - (3 if R > 8 else 0)
            q = synth.make_query(9000 + lq, lq)
            if lq not in want:
                want[lq] = oracle.search32(q, codes, off, tab, go, ge)
            os.environ["SWGPU_K1S_PLAN"] = f"{g},{R}"
            got, st = db.search(q, tab, go, ge)
            assert (st["lanes_per_pair"], st["tile_rows"], st["passes"]) == (g, R, 1), (g, R, st)
            assert np.array_equal(got, want[lq]), (g, R, int((got != want[lq]).sum()))
    db.close()


def test_generic_kernel_equals_the_gap_specialisations(oracle, matrices, plan_env):
    """SWGPU_K1S_GENERIC=1: the -10/-1 search through the generic kernel (penalties in registers)."""
    import swgpu
    from swgpu import synth
    tab = matrices["BLOSUM62"]
    codes, off, _ = small_db(synth, seed=75, n=600)
    db = swgpu.Database(codes, off)
    for lq in (144, 375, 1000, 2005):
        q = synth.make_query(9400 + lq, lq)
        os.environ.pop("SWGPU_K1S_GENERIC", None)
        a, _ = db.search(q, tab, -10, -1)
        os.environ["SWGPU_K1S_GENERIC"] = "1"
        b, _ = db.search(q, tab, -10, -1)
        assert np.array_equal(a, b) and np.array_equal(a, oracle.search32(q, codes, off, tab, -10, -1)), lq
    db.close()


@pytest.mark.parametrize("go,ge", [(-12, -2), (-10, -1), (-11, -1)])
def test_every_multi_stripe_instantiation_small_database(oracle, matrices, plan_env, go, ge):
    import swgpu
    from swgpu import synth
    tab = matrices["PAM250"]
    codes, off, _ = small_db(synth, seed=72, n=700)
    db = swgpu.Database(codes, off)
    for R in MULTI_R:
        for npass in (2, 3):
            lq = 32 * R * npass - 5
            q = synth.make_query(9100 + lq, lq)
            want = oracle.search32(q, codes, off, tab, go, ge)
            os.environ["SWGPU_K1S_PLAN"] = f"32,{R}"
            got, st = db.search(q, tab, go, ge)
            assert (st["lanes_per_pair"], st["tile_rows"], st["passes"]) == (32, R, npass), (R, st)
            assert np.array_equal(got, want), (R, npass, int((got != want).sum()))
    db.close()


@pytest.mark.parametrize("go,ge", GAPS)
def test_every_multi_stripe_instantiation_ordinary_items(oracle, matrices, plan_env, go, ge):
    import swgpu
    from swgpu import synth
    tab = matrices["BLOSUM62"]
    codes, off = synth.make_db(73, 90000, hi=12000, force_max=12000, sort=False)
    n = len(off) - 1
    db = swgpu.Database(codes, off)
    sample = np.concatenate([np.arange(0, n, 211), np.argsort(off[1:] - off[:-1])[-6:]])
    parts = [codes[off[t]:off[t + 1]] for t in sample]
    so = np.zeros(len(sample) + 1, dtype=np.int64); np.cumsum([len(p) for p in parts], out=so[1:])
    sc = np.concatenate(parts)
    for R in MULTI_R:
        lq = 64 * R - 7
        q = synth.make_query(9200 + lq, lq)
        os.environ["SWGPU_K1S_PLAN"] = f"32,{R}"
        got, st = db.search(q, tab, go, ge)
        assert (st["tile_rows"], st["passes"]) == (R, 2), (R, st)
        want = oracle.search32(q, sc, so, tab, go, ge)
        assert np.array_equal(got[sample], want), (R, int((got[sample] != want).sum()))
    db.close()


@pytest.mark.parametrize("stream_cols,no_taper", [(64, 0), (300, 1), (1536, 0), (2304, 1), (0, 0), (8192, 0)])
def test_packing_options_do_not_change_scores(oracle, matrices, stream_cols, no_taper):
    import swgpu
    from swgpu import synth
    tab = matrices["BLOSUM45"]
    codes, off, _ = small_db(synth, seed=74, n=900)
    db = swgpu.Database(codes, off, stream_cols=stream_cols, no_taper=no_taper)
    for lq in (57, 200, 333, 729, 1000, 1500):
        q = synth.make_query(9300 + lq, lq)
        got, _ = db.search(q, tab, -10, -1)
        want = oracle.search32(q, codes, off, tab, -10, -1)
        assert np.array_equal(got, want), (stream_cols, no_taper, lq, int((got != want).sum()))
    hits = db.topk(10)
    for h in hits[:3]:
        e = h["entry"]
        s, c = oracle.end_start(q, codes[off[e]:off[e + 1]], tab, -10, -1)
        assert (h["score"], h["q_start"], h["q_end"], h["t_start"], h["t_end"]) == (s, *c)
    db.close()
