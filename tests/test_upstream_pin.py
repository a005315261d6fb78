"""CPU suite: coordinates and alignment paths pinned against something that is NOT this repo's own restatement --
the upstream seq-align ELF the reference's differential test drives (/root/reference/test/smith_waterman,
test/tests.py:58-73, test/run_tests.sh:1).  tests/golden/make_golden_r2.py froze its `hit 0.0` (score, [pos; len]
of both sequences, the printed alignment) for 208 planted pairs under {PAM250, default gaps -2/-1} and
{BLOSUM62, -10/-1}; its source is not in the reference tree, so its tie-break cannot be audited:

  * the SCORE is gated on every pair (oracle == upstream);
  * the END cell is gated wherever only one cell attains the best score (then every tie-break must agree);
  * the START cell is gated wherever, in addition, the reversed problem from that end cell has a unique optimum;
  * the PATH (CIGAR) is gated to be co-optimal: it must score exactly the hit's score under the reference's gap
    convention, and it is compared op-for-op with upstream's where both coordinates agree (reported, and required
    to agree in >= 90 % of the BLOSUM62 -10/-1 pairs; with PAM250 and the cheap default gaps -2/-1 most optima have
    many equal-score paths and the two tie-break rules pick different ones -- there only co-optimality is gated).

The GPU twin (tests/test_gpu_upstream.py, -m gpu) replays the same pairs through swgpu_topk / swgpu_traceback."""
import os

import numpy as np

from conftest import SCORING, golden


def load_pairs():
    from swgpu import synth
    g = golden("upstream_hits.npz")
    tabs, gaps = [], []
    for s in g["settings"]:
        name, gp = str(s).split(":")
        tabs.append(synth.load_matrix(os.path.join(SCORING, name + ".txt"))[0])
        gaps.append((-2, -1) if gp == "default" else tuple(int(x) for x in gp.split(",")))
    enc = lambda s: np.array([ord(c) - 64 for c in str(s)], dtype=np.int8)
    out = []
    for i in range(len(g["score"])):
        si = int(g["setting"][i])
        up = dict(score=int(g["score"][i]), q_start=int(g["q_pos"][i]), q_end=int(g["q_pos"][i]) + int(g["q_len"][i]) - 1,
                  t_start=int(g["t_pos"][i]), t_end=int(g["t_pos"][i]) + int(g["t_len"][i]) - 1,
                  q_aln=str(g["q_aln"][i]), t_aln=str(g["t_aln"][i]))
        out.append((enc(g["q"][i]), enc(g["t"][i]), tabs[si], gaps[si], up))
    return out


def aln_to_cigar(q_aln, t_aln):
    """upstream's two printed rows -> CIGAR in this repo's convention (I = query residue against a gap)."""
    ops = []
    for a, b in zip(q_aln, t_aln):
        op = "D" if a == "-" else "I" if b == "-" else "M"
        if ops and ops[-1][1] == op:
            ops[-1][0] += 1
        else:
            ops.append([1, op])
    return "".join(f"{n}{o}" for n, o in ops)


def test_scores_coordinates_and_paths_against_upstream_binary(oracle):
    pairs = load_pairs()
    assert len(pairs) >= 200
    n_end_gated = n_start_gated = 0
    cmp_by_gap, same_by_gap = {}, {}
    for q, t, tab, (go, ge), up in pairs:
        s, (qs, qe, ts, te) = oracle.end_start(q, t, tab, go, ge)
        assert s == up["score"], (s, up)
        assert oracle.score32(q, t, tab, go, ge) == up["score"]
        end_unique = oracle.count_best(q, t, tab, go, ge) == 1
        if end_unique:
            n_end_gated += 1
            assert (qe, te) == (up["q_end"], up["t_end"]), ((qs, qe, ts, te), up)
            start_unique = oracle.count_best(q[:qe + 1][::-1], t[:te + 1][::-1], tab, go, ge) == 1
            if start_unique:
                n_start_gated += 1
                assert (qs, ts) == (up["q_start"], up["t_start"]), ((qs, qe, ts, te), up)
        # the path this repo reports for its own rectangle: co-optimal by an independent re-score
        cg, sc = oracle.traceback(q[qs:qe + 1], t[ts:te + 1], tab, go, ge)
        assert sc == s and oracle.cigar_score(cg, q[qs:qe + 1], t[ts:te + 1], tab, go, ge) == s
        # upstream's own path re-scored under the reference's gap convention (open + N * extend)
        ucg = aln_to_cigar(up["q_aln"], up["t_aln"])
        uq, ut = q[up["q_start"]:up["q_end"] + 1], t[up["t_start"]:up["t_end"] + 1]
        assert oracle.cigar_score(ucg, uq, ut, tab, go, ge) == s
        if (qs, qe, ts, te) == (up["q_start"], up["q_end"], up["t_start"], up["t_end"]):
            cmp_by_gap[go] = cmp_by_gap.get(go, 0) + 1
            same_by_gap[go] = same_by_gap.get(go, 0) + (cg == ucg)
    assert n_end_gated >= 120 and n_start_gated >= 100, (n_end_gated, n_start_gated)
    assert cmp_by_gap[-10] >= 90 and same_by_gap[-10] >= 0.9 * cmp_by_gap[-10], (cmp_by_gap, same_by_gap)
    print(f"upstream pin: {len(pairs)} scores, {n_end_gated} end cells, {n_start_gated} start cells gated; paths identical op for "
          f"op: {same_by_gap[-10]}/{cmp_by_gap[-10]} (BLOSUM62 -10/-1), {same_by_gap[-2]}/{cmp_by_gap[-2]} (PAM250 -2/-1)")
