"""CPU suite: pins the oracle (oracle/sw_oracle.c) against the golden vectors produced by the
UNMODIFIED reference (tests/golden/make_golden.py) and, where oracle/_ref exists, against the
reference library itself.  No GPU needed."""
import numpy as np
import pytest

from conftest import golden

GAPS = [(-10, -1), (-2, -1), (0, -1), (-11, -1), (-5, -3), (-1, -1)]


@pytest.mark.parametrize("name", ["BLOSUM62", "BLOSUM45", "PAM250"])
def test_int32_oracle_matches_reference_golden(oracle, name):
    g = golden(f"ref16_{name}.npz")
    codes, off, q, tab = g["codes"], g["off"], g["query"], g["table"]
    checked = 0
    for lq in (1, 7, 33, 144, 360, 700):
        for go, ge in GAPS:
            want = g[f"s_{lq}_{-go}_{-ge}"].astype(np.int32)
            got = oracle.search32(q[:lq], codes, off, tab, go, ge)
            assert np.array_equal(got, want), (name, lq, go, ge)
            checked += want.size
    assert checked == 6 * 6 * 272


def test_table_loader_matches_golden_tables(matrices):
    for name, tab in matrices.items():
        assert np.array_equal(tab, golden(f"ref16_{name}.npz")["table"])
        assert tab.shape == (32, 32) and tab[31, 31] == 1          # '*' x '*' = +1 in every shipped matrix


def test_wrapping_int16_restatement_matches_reference_on_overflow(oracle):
    g = golden("ref16_overflow.npz")
    codes, off, q, tab = g["codes"], g["off"], g["query"], g["table"]
    ref16, true32 = g["ref16"], g["true32"]
    got32 = oracle.search32(q, codes, off, tab, -10, -1)
    assert np.array_equal(got32, true32)
    assert true32.min() < 32767 < true32.max()                      # the set straddles the int16 limit
    # the reference is exact iff the true score fits; above that it prints wrapped garbage
    fits = true32 <= 32767
    assert np.array_equal(ref16[fits].astype(np.int32), true32[fits])
    assert np.all(ref16[~fits].astype(np.int32) != true32[~fits])
    # the wrap-exact restatement reproduces even the garbage (batch of 16, padded to the first member)
    n = off.size - 1
    for k in range(n):
        padded = int(off[(k // 16) * 16 + 1] - off[(k // 16) * 16])
        w = oracle.score16_wrap(q, codes[off[k]:off[k + 1]], padded, tab, -10, -1)
        assert w == int(ref16[k]), k
    # overflow predicate: a 16-bit result <= 32767 - max(matrix) is always exact
    safe = ref16.astype(np.int32) <= 32767 - int(tab.max())
    assert np.array_equal(ref16[safe].astype(np.int32), true32[safe])


def test_live_reference_library_agrees(oracle, matrices):
    if not oracle.have_ref():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    from swgpu import synth
    codes, off = synth.make_db(11, 160, median=80.0, hi=500)
    q = synth.make_query(12, 123)
    for name, tab in matrices.items():
        for go, ge in ((-10, -1), (-3, -2)):
            ref, _ = oracle.ref_search16(q, codes, off, tab, go, ge, threads=2)
            assert np.array_equal(ref.astype(np.int32), oracle.search32(q, codes, off, tab, go, ge)), name


def test_letters_to_index(oracle):
    enc = golden("encode.npz")["enc"]
    for c in range(256):
        assert oracle.oracle().swo_letters_to_index(c) == enc[c], c
    assert enc[ord("A")] == 1 and enc[ord("z")] == 26 and enc[ord("*")] == 31 and enc[ord("-")] == -1


def test_cli_golden_scores(oracle, matrices):
    g = golden("cli_small.npz")
    got = oracle.search32(g["query"], g["codes"], g["off"], matrices["BLOSUM62"], -10, -1)
    assert np.array_equal(got, g["scores"])
    out = str(g["stdout_plain"])
    assert out.startswith("Query File=") and "Total Entries: 32" in out and "Entry #31:" in out


def test_coordinates_rule(oracle, matrices):
    """End/start rule of SURVEY.md A.7 (PARITY UNPINNED: the reference keeps only a max)."""
    from swgpu import synth
    tab = matrices["BLOSUM62"]
    g = synth.rng(5)
    q = synth.make_query(6, 150)
    for trial in range(40):
        t = synth.residues(g, int(g.integers(5, 300)))
        if trial % 3 == 0:                       # plant a related segment
            a, L = int(g.integers(0, 100)), int(g.integers(10, 50))
            t = np.concatenate([t[:20], q[a:a + L], t[20:]])
        s, (qs, qe, ts, te) = oracle.end_start(q, t, tab, -10, -1)
        assert s == oracle.score32(q, t, tab, -10, -1)
        if s == 0:
            assert (qs, qe, ts, te) == (-1, -1, -1, -1)
            continue
        assert 0 <= qs <= qe < q.size and 0 <= ts <= te < t.size
        # the reported box contains an alignment of the full score
        assert oracle.score32(q[qs:qe + 1], t[ts:te + 1], tab, -10, -1) == s
        # end = first maximum in scan order: no earlier target prefix reaches the score
        assert oracle.score32(q, t[:te], tab, -10, -1) < s
        assert oracle.score32(q[:qe], t[:te + 1], tab, -10, -1) < s


def test_empty_inputs(oracle, matrices):
    tab = matrices["PAM250"]
    assert oracle.score32(np.zeros(0, np.int8), np.array([1, 2, 3], np.int8), tab, -10, -1) == 0
    assert oracle.score32(np.array([1, 2, 3], np.int8), np.zeros(0, np.int8), tab, -10, -1) == 0
