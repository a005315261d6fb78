"""Positive gap scores.  The reference accepts any --gapopen / --gapextend (src/alignment_cmdline.c:255-267) and runs its
three-state recurrence on them (src/alignment.c:124-161); the engine's two-gap-state kernels are only score-equivalent
for non-positive values, so a positive gap score routes to k3_three (csrc/sw_k3s.cuh).  Goldens: tests/golden/
ref16_posgap.npz, frozen from the unmodified reference (make_golden_r2.py posgap):
  * `s_*`  batches of 16 EQUAL-length targets (no '*' padding rows): the per-target truth -- oracle and GPU must match;
  * `r_*`  ragged batches: with positive gap scores the reference's padding rows change the result (E / F keep growing
           across them), so these pin the PADDED computation, which swgpu_fill_batch16 / alignment_fill_matrices reproduce;
  * the reference CLI's stdout for --gapopen 1 --gapextend 1."""
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT, SCORING, golden

POS_GAPS = [(1, -3), (2, -1), (0, 1), (1, 1), (-5, 2), (3, 0)]
LQS = (40, 130, 300)
BIN = os.path.join(ROOT, "bin", "smith_waterman")


def test_oracle_three_state_equals_reference_with_positive_gaps(oracle):
    g = golden("ref16_posgap.npz")
    q, tab = g["query"], g["table"]
    for lq in LQS:
        for go, ge in POS_GAPS:
            got = oracle.search32(q[:lq], g["codes"], g["off"], tab, go, ge)
            assert np.array_equal(got, g[f"s_{lq}_{go}_{ge}"].astype(np.int32)), (lq, go, ge)
    # the padded arithmetic (what alignment_fill_matrices really computes for a ragged batch)
    roff, rcodes = g["roff"], g["rcodes"]
    for go, ge in POS_GAPS:
        want = g[f"r_130_{go}_{ge}"]
        for t in range(0, 48, 5):
            plen = int(roff[t // 16 * 16 + 1] - roff[t // 16 * 16])
            assert oracle.score16_wrap(q[:130], rcodes[roff[t]:roff[t + 1]], plen, tab, go, ge) == int(want[t]), (go, ge, t)


@pytest.mark.gpu
def test_gpu_scores_with_positive_gaps_equal_reference_goldens(oracle):
    import swgpu
    g = golden("ref16_posgap.npz")
    q, tab = g["query"], g["table"]
    db = swgpu.Database(g["codes"], g["off"])
    for lq in LQS:
        for go, ge in POS_GAPS:
            got, st = db.search(q[:lq], tab, go, ge)
            assert np.array_equal(got, g[f"s_{lq}_{go}_{ge}"].astype(np.int32)), (lq, go, ge)
            assert st["tile_rows"] == 0 and st["n_rescored"] == len(got)             # no 16-bit pass: every target through the 32-bit kernel
    # the pipelined entry point too; top-k with coordinates is refused (its coordinate passes use the two-gap-state form)
    scores, _, _, _ = db.search_many([q[:130], q[:300]], tab, 1, 1)
    assert np.array_equal(scores[1], g["s_300_1_1"].astype(np.int32))
    with pytest.raises(swgpu.SwgpuError) as e:
        db.search_many([q[:130]], tab, 1, 1, k=5)
    assert "positive gap" in str(e.value)
    got, _ = db.search(q[:130], tab, -10, -1)                                         # the handle is still usable afterwards
    assert np.array_equal(got, oracle.search32(q[:130], g["codes"], g["off"], tab, -10, -1))
    db.close()


@pytest.mark.gpu
def test_gpu_positive_gaps_long_targets_many_stripes_and_ragged_lengths(oracle, matrices):
    """beyond the goldens' sizes: a 5 000-aa target (carry rows through global memory), 700-row query (3 stripes), empty
    targets, every residue code incl. '*' and letters outside the matrix -- against the three-state int32 oracle"""
    import swgpu
    from swgpu import synth
    g = synth.rng(808)
    lens = np.concatenate([[5000, 0, 1, 33], g.integers(1, 400, size=200)])
    off = np.zeros(len(lens) + 1, dtype=np.int64); np.cumsum(lens, out=off[1:])
    codes = g.integers(1, 32, size=int(off[-1])).astype(np.int8)
    q = g.integers(1, 32, size=700).astype(np.int8)
    codes[off[0] + 100: off[0] + 700] = q[:600]
    db = swgpu.Database(codes, off)
    for lq, (go, ge) in [(700, (1, -2)), (100, (0, 1)), (129, (2, 2)), (700, (-3, 1))]:
        got, _ = db.search(q[:lq], matrices["PAM250"], go, ge)
        want = oracle.search32(q[:lq], codes, off, matrices["PAM250"], go, ge)
        bad = np.nonzero(got != want)[0]
        assert bad.size == 0, (lq, go, ge, bad[:6], got[bad[:6]], want[bad[:6]])
    db.close()


@pytest.mark.gpu
def test_fill_batch16_reproduces_the_reference_padded_computation_with_positive_gaps():
    import swgpu
    g = golden("ref16_posgap.npz")
    q, tab, roff, rcodes = g["query"], g["table"], g["roff"], g["rcodes"]
    for b in range(3):
        plen = int(roff[b * 16 + 1] - roff[b * 16])
        batch = np.full((plen, 16), 31, dtype=np.int8)
        for l in range(16):
            t = rcodes[roff[b * 16 + l]:roff[b * 16 + l + 1]]
            batch[:len(t), l] = t
        for lq in LQS:
            for go, ge in POS_GAPS:
                got = swgpu.fill_batch16(q[:lq], batch.reshape(-1), plen, 16, tab, go, ge)
                assert list(got) == list(g[f"r_{lq}_{go}_{ge}"][b * 16:(b + 1) * 16].astype(np.int32)), (b, lq, go, ge)


@pytest.mark.gpu
def test_cli_stdout_with_positive_gaps_equals_reference_binary(tmp_path):
    g = golden("ref16_posgap.npz")
    q = tmp_path / "q.fa"; q.write_text(str(g["cli_query_fasta"]))
    d = tmp_path / "db.fa"; d.write_text(str(g["cli_db_fasta"]))
    r = subprocess.run([BIN, "--substitution_matrix", os.path.join(SCORING, "BLOSUM62.txt"), "--gapopen", "1", "--gapextend", "1",
                        "--files", str(q), str(d)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    norm = lambda s: re.sub(r"Total Time: [0-9.]+", "Total Time: T", s)
    assert norm(r.stdout.replace(str(tmp_path), "TMP")) == norm(str(g["cli_stdout"]))
