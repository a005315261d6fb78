"""GPU suite (-m gpu): the CUDA path, called through the C-ABI (libswgpu.so), against the CPU oracle
and the committed reference golden vectors.  Integer work: the bar is bit-exact."""
import numpy as np
import pytest

from conftest import golden

pytestmark = pytest.mark.gpu
GAPS = [(-10, -1), (-2, -1), (0, -1), (-11, -1), (-5, -3), (-1, -1)]


@pytest.fixture(scope="module")
def swgpu_mod():
    import swgpu
    assert swgpu.lib().swgpu_device_count() >= 1, "GPU suite needs a CUDA device; there is no CPU fallback"
    return swgpu


def check(swgpu, oracle, q, codes, off, tab, go, ge, **kw):
    db = swgpu.Database(codes, off, **kw)
    got, st = db.search(q, tab, go, ge)
    db.close()
    want = oracle.search32(q, codes, off, tab, go, ge)
    bad = np.nonzero(got != want)[0]
    assert bad.size == 0, (bad[:8], got[bad[:8]], want[bad[:8]])
    return got, st


@pytest.mark.parametrize("name", ["BLOSUM62", "BLOSUM45", "PAM250"])
def test_reference_golden_vectors(swgpu_mod, name):
    """Scores produced by the unmodified reference (tests/golden/make_golden.py)."""
    g = golden(f"ref16_{name}.npz")
    codes, off, q, tab = g["codes"], g["off"], g["query"], g["table"]
    for long_thr in (0, 200):                       # 16-bit only, and targets > 200 aa re-scored by the 32-bit kernel
        db = swgpu_mod.Database(codes, off, long_threshold=long_thr)
        for lq in (1, 7, 33, 144, 360, 700):
            for go, ge in GAPS:
                got, _ = db.search(q[:lq], tab, go, ge)
                assert np.array_equal(got, g[f"s_{lq}_{-go}_{-ge}"].astype(np.int32)), (name, lq, go, ge, long_thr)
        db.close()


def test_reference_cli_golden(swgpu_mod):
    g = golden("cli_small.npz")
    from swgpu import synth
    import os
    from conftest import SCORING
    tab, _ = synth.load_matrix(os.path.join(SCORING, "BLOSUM62.txt"))
    db = swgpu_mod.Database(g["codes"], g["off"])
    got, _ = db.search(g["query"], tab, -10, -1)
    assert np.array_equal(got, g["scores"])


def test_config1_full(swgpu_mod, oracle, matrices):
    """C1: 360-aa query x 10 000 proteins, BLOSUM62, -10/-1 -- every score against the scalar oracle."""
    from swgpu import synth
    q, codes, off = synth.config1()
    got, st = check(swgpu_mod, oracle, q, codes, off, matrices["BLOSUM62"], -10, -1)
    assert st["cells"] == 360 * int(off[-1]) and st["kernel_launches"] >= 3
    # file order is preserved for an UNSORTED database too (the engine sorts internally)
    perm = synth.rng(3).permutation(len(off) - 1)
    lens = (off[1:] - off[:-1])[perm]
    off2 = np.zeros_like(off); np.cumsum(lens, out=off2[1:])
    codes2 = np.concatenate([codes[off[p]:off[p + 1]] for p in perm])
    db = swgpu_mod.Database(codes2, off2)
    got2, _ = db.search(q, matrices["BLOSUM62"], -10, -1)
    assert np.array_equal(got2, got[perm])


@pytest.mark.parametrize("name", ["BLOSUM62", "BLOSUM45", "PAM250"])
@pytest.mark.parametrize("lq", [1, 2, 15, 16, 17, 31, 32, 33, 63, 64, 65, 100, 257, 1000])
def test_query_lengths_and_tile_edges(swgpu_mod, oracle, matrices, name, lq):
    from swgpu import synth
    codes, off = synth.make_db(100 + lq, 700, median=90.0, lo=1, hi=1200)
    q = synth.make_query(200 + lq, lq)
    check(swgpu_mod, oracle, q, codes, off, matrices[name], -10, -1)


@pytest.mark.parametrize("go,ge", GAPS + [(-40, -4), (0, 0), (-1, 0), (-100, -100)])
def test_gap_settings(swgpu_mod, oracle, matrices, go, ge):
    from swgpu import synth
    codes, off = synth.make_db(31, 500, median=120.0, lo=1, hi=900)
    q = synth.make_query(32, 222)
    # related targets so gaps matter
    g = synth.rng(33)
    for k in range(0, 500, 3):
        L = int(off[k + 1] - off[k])
        src = np.resize(q, L).copy()
        mut = g.random(L) < 0.2
        src[mut] = synth.residues(g, int(mut.sum()))
        cut = int(g.integers(0, max(L - 4, 1)))
        src = np.concatenate([src[:cut], src[cut + 3:], src[:3]])[:L]
        codes[off[k]:off[k + 1]] = src
    check(swgpu_mod, oracle, q, codes, off, matrices["BLOSUM62"], go, ge)


def test_ragged_and_degenerate_databases(swgpu_mod, oracle, matrices):
    from swgpu import synth
    tab = matrices["BLOSUM62"]
    q = synth.make_query(1, 77)
    # empty database
    db = swgpu_mod.Database(np.zeros(0, np.int8), np.zeros(1, np.int64))
    got, _ = db.search(q, tab, -10, -1)
    assert got.size == 0
    # one target; 63 / 64 / 65 targets; empty targets in the middle; N not a multiple of 16 or 64
    for n in (1, 2, 15, 63, 64, 65, 129, 1000):
        g = synth.rng(n)
        lens = g.integers(0, 300, size=n)
        lens[g.integers(0, n)] = 0
        off = np.zeros(n + 1, np.int64); np.cumsum(lens, out=off[1:])
        codes = synth.residues(g, off[-1])
        check(swgpu_mod, oracle, q, codes, off, tab, -10, -1)
    # empty query -> all zeros
    codes, off = synth.make_db(2, 100, median=50.0, lo=1, hi=100)
    db = swgpu_mod.Database(codes, off)
    got, _ = db.search(np.zeros(0, np.int8), tab, -10, -1)
    assert np.array_equal(got, np.zeros(100, np.int32))
    # every residue code the encoder can produce, including '*' (31) and letters outside the matrix (J O U)
    g = synth.rng(9)
    codes = g.integers(1, 32, size=20000).astype(np.int8)
    off = np.arange(0, 20001, 100, dtype=np.int64)
    qq = g.integers(1, 32, size=150).astype(np.int8)
    check(swgpu_mod, oracle, qq, codes, off, tab, -10, -1)


def test_overflow_golden_and_rescoring(swgpu_mod, oracle):
    """C3 core: PAM250 targets whose true scores straddle 32 767 (reference output recorded beside)."""
    g = golden("ref16_overflow.npz")
    codes, off, q, tab = g["codes"], g["off"], g["query"], g["table"]
    db = swgpu_mod.Database(codes, off, long_threshold=100000)     # force all through the 16-bit kernel first
    got, st = db.search(q, tab, -10, -1)
    assert np.array_equal(got, g["true32"])
    fits = g["true32"] <= 32767
    assert np.array_equal(got[fits], g["ref16"][fits].astype(np.int32))          # == reference where it is exact
    # the engine's 16-bit path is UNSIGNED with a bias B = 128 + |oe| + |ge|: it stays exact up to
    # 65535 - 129 - B, so none of these (true scores <= 34 752) needs the 32-bit pass
    assert st["n_rescored"] == 0


def test_config3_overflow_stress_scaled(swgpu_mod, oracle, matrices):
    """C3 (scaled to 384 targets so the scalar oracle finishes in seconds): far-overflow, boundary,
    mutated and background targets, PAM250."""
    from swgpu import synth
    tab = matrices["PAM250"]
    g = synth.rng(44)
    lc = np.array([23, 3, 25, 6], dtype=np.int8)
    q = lc[g.choice(4, size=6000, p=[0.4, 0.2, 0.2, 0.2])]
    parts = []
    for _ in range(96):                                   # (A) far overflow
        L = int(g.integers(3000, 6001)); s = int(g.integers(0, 6000 - L + 1)); parts.append(q[s:s + L])
    for _ in range(96):                                   # (B) boundary
        L = int(g.integers(2300, 2701)); s = int(g.integers(0, 6000 - L + 1)); parts.append(q[s:s + L])
    for k in range(96):                                   # (C) mutated copies
        src = parts[k * 2].copy()
        mut = g.random(src.size) < g.uniform(0.01, 0.05)
        src[mut] = synth.residues(g, int(mut.sum()))
        cut = int(g.integers(10, src.size - 10))
        parts.append(np.concatenate([src[:cut], src[cut + 2:]]))
    bg, bo = synth.make_db(45, 96)
    parts += [bg[bo[i]:bo[i + 1]] for i in range(96)]     # (D) background
    order = g.permutation(len(parts))
    parts = [parts[i] for i in order]
    off = np.zeros(len(parts) + 1, np.int64); np.cumsum([p.size for p in parts], out=off[1:])
    codes = np.concatenate(parts)
    want = oracle.search32(q, codes, off, tab, -10, -1)
    assert int((want > 32767).sum()) >= 96                    # beyond the reference's int16 range
    n_over = int((want > 65535 - 129 - (128 + 11 + 1)).sum())      # beyond the engine's biased-u16 range
    assert n_over >= 10
    for thr in (0, 100000):
        db = swgpu_mod.Database(codes, off, long_threshold=thr)
        got, st = db.search(q, tab, -10, -1)
        assert np.array_equal(got, want)
        if thr == 100000:
            assert st["n_rescored"] >= n_over
        db.close()


def test_config4_long_sequences_scaled(swgpu_mod, oracle, matrices):
    """C4 (scaled): 35 213-aa query against targets up to 35 213 aa, incl. a 90 %-identity copy that
    overflows 16 bits."""
    from swgpu import synth
    tab = matrices["BLOSUM62"]
    g = synth.rng(45)
    q = synth.make_query(4500, 35213)
    lens = [35213, 35213, 20000, 9000] + list(synth.lognormal_lengths(g, 60))
    parts = [synth.residues(g, L) for L in lens]
    cp = q.copy(); mut = g.random(cp.size) < 0.10; cp[mut] = synth.residues(g, int(mut.sum()))
    parts[1] = cp
    off = np.zeros(len(parts) + 1, np.int64); np.cumsum([p.size for p in parts], out=off[1:])
    codes = np.concatenate(parts)
    got, st = check(swgpu_mod, oracle, q, codes, off, tab, -10, -1)
    assert got[1] > 65535 and st["n_rescored"] >= 1


def test_32bit_kernels_on_everything(swgpu_mod, oracle, matrices):
    """long_threshold=1 sends every target through the 32-bit kernels: k3_intra (< 4096 aa) and the
    wavefront kernel k3x (>= 4096 aa, 22 stripes for the 2 700-aa query)."""
    from swgpu import synth
    g = synth.rng(77)
    lens = [9000, 6500, 4096, 4095, 5000] + list(synth.lognormal_lengths(g, 300))
    parts = [synth.residues(g, int(L)) for L in lens]
    q = synth.make_query(78, 2700)
    parts[1][1000:3000] = q[300:2300]
    off = np.zeros(len(parts) + 1, np.int64); np.cumsum([p.size for p in parts], out=off[1:])
    codes = np.concatenate(parts)
    for lq in (2700, 100, 129):
        got, st = check(swgpu_mod, oracle, q[:lq], codes, off, matrices["BLOSUM45"], -12, -2, long_threshold=1)
        assert st["n_rescored"] == len(parts) and st["cells_rescore"] == st["cells"] == lq * int(off[-1])   # summed on the device
    # huge gap penalties: the 16-bit bias does not fit -> the 32-bit kernels score the whole database directly
    got, st = check(swgpu_mod, oracle, q[:300], codes, off, matrices["BLOSUM62"], -20000, -1)
    assert st["tile_rows"] == 0


def test_resident_search_and_d2h(swgpu_mod, oracle, matrices):
    from swgpu import synth
    q, codes, off = synth.config1()
    db = swgpu_mod.Database(codes[:off[2000]], off[:2001])
    st = db.search_resident(q, matrices["BLOSUM45"], -10, -1)
    assert np.array_equal(db.scores(), oracle.search32(q, codes[:off[2000]], off[:2001], matrices["BLOSUM45"], -10, -1))
    assert st["ms_total"] > 0


def test_fill_batch16_reference_layout(swgpu_mod, oracle, matrices):
    """GPU twin of alignment_fill_matrices(aligner_t*): the reference's own [pos][lane] x 16 layout with
    '*' padding; scores must equal the reference's including the padded rows."""
    from swgpu import synth
    tab = matrices["BLOSUM62"]
    g = synth.rng(16)
    lens = np.sort(g.integers(5, 120, size=16))[::-1]
    plen = int(lens[0])
    batch = np.full((plen, 16), 31, dtype=np.int8)
    tg = [synth.residues(g, int(L)) for L in lens]
    for l in range(16):
        batch[:lens[l], l] = tg[l]
    q = synth.make_query(17, 90)
    got = swgpu_mod.fill_batch16(q, batch.reshape(-1), plen, 16, tab, -10, -1)
    want = [oracle.score16_wrap(q, tg[l], plen, tab, -10, -1) for l in range(16)]
    assert list(got) == want


def test_config4_full_size(swgpu_mod, oracle, matrices):
    """C4 at full size: 35 213-aa query x 1 024 targets up to 35 213 aa, 62 G cells.  The scalar
    oracle needs ~2 s per titin-length target, so it checks a sample; every target is cross-checked between the two
    independent GPU paths (packed 16-bit K1S with X items + rescoring vs. the 32-bit kernels on everything)."""
    from swgpu import synth
    tab = matrices["BLOSUM62"]
    q, codes, off = synth.config4()
    db = swgpu_mod.Database(codes, off)
    got, st = db.search(q, tab, -10, -1)
    db.close()
    assert st["passes"] > 1 and st["n_rescored"] >= 1 and got[1] > 65535
    for i in (0, 1, 15, 40, 63, 64, 500, 1023):
        assert got[i] == oracle.score32(q, codes[off[i]:off[i + 1]], tab, -10, -1), i
    db32 = swgpu_mod.Database(codes, off, long_threshold=1)
    got32, st32 = db32.search(q, tab, -10, -1)
    db32.close()
    assert st32["n_rescored"] == len(off) - 1
    assert np.array_equal(got, got32)
