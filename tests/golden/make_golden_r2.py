"""Round-2 golden vectors (run in the dev container only: needs /root/reference and oracle/_ref):

    python tests/golden/make_golden_r2.py [c2] [c3] [c4] [upstream]

  c2_sample.npz      the BENCHMARK configuration (C2, seed 43, 570 000 targets / 205 M residues): int16 scores of the
                     unmodified reference kernel (oracle/_ref) for every 512-th batch of 16 consecutive targets and
                     the query lengths 144 / 375 / 1000 / 2005 / 5478.  The GPU box regenerates the database from the
                     seed and replays these (tests/test_gpu_c2.py).
  c3_full.npz        C3 at SURVEY.md 8(d)'s size (4 096 targets, Lq = 6 000, PAM250): the reference's int16 output
                     (wrapped where the true score exceeds 32 767) and the int32 scalar oracle's true scores.
  c4_sample.npz      C4 (35 213-aa query x 1 024 targets): int32 scalar-oracle scores of 120 targets incl. all 16
                     titin-length ones and the 90 %-identity copy.
  ref16_posgap.npz   positive gap scores (the reference's three-state recurrence): the unmodified kernel's scores for 272
                     short targets x 3 query lengths x 6 gap settings, and the reference CLI's stdout for --gapopen 1 --gapextend 1.
  upstream_hits.npz  the upstream seq-align ELF the reference's own differential test drives
                     (/root/reference/test/smith_waterman, test/tests.py:58-73): `hit 0.0` of ~200 planted pairs under
                     {PAM250, default gaps -2/-1} and {BLOSUM62, -10/-1}: score, [pos; len] of both sequences and
                     the printed alignment.  The only evidence for coordinates / paths that is not this repo's own.
"""
import os, re, subprocess, sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "seq-align-gpu_b200"))
from swgpu import synth            # noqa: E402
from oracle import oracle as O     # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
UPSTREAM = "/root/reference/test/smith_waterman"
C2_LQ = (144, 375, 1000, 2005, 5478)
THREADS = len(os.sched_getaffinity(0))


def sample_groups(off, stride):
    n = len(off) - 1
    groups = np.arange(stride // 2, n // 16, stride)
    return (groups[:, None] * 16 + np.arange(16)[None, :]).reshape(-1)


def take(codes, off, idx):
    lens = (off[1:] - off[:-1])[idx]
    o = np.zeros(len(idx) + 1, dtype=np.int64); np.cumsum(lens, out=o[1:])
    return np.concatenate([codes[off[i]:off[i + 1]] for i in idx]), o


def gen_c2():
    tab, _ = synth.load_matrix(os.path.join(ROOT, "scoring", "BLOSUM62.txt"))
    codes, off = synth.config2()
    idx = sample_groups(off, 512)
    c, o = take(codes, off, idx)
    res = {}
    for lq in C2_LQ:
        q = synth.make_query(100 + synth.C2_QUERY_LENGTHS.index(lq), lq)
        s, _ = O.ref_search16(q, c, o, tab, -10, -1, threads=THREADS)
        res[f"ref16_{lq}"] = s
        print("c2 Lq", lq, "max", int(s.max()))
    np.savez_compressed(os.path.join(OUT, "c2_sample.npz"), idx=idx, n_targets=len(off) - 1, n_residues=int(off[-1]),
                        sample_residues=int(o[-1]), **res)


def gen_c3():
    tab, _ = synth.load_matrix(os.path.join(ROOT, "scoring", "PAM250.txt"))
    q, codes, off = synth.config3()
    s16, _ = O.ref_search16(q, codes, off, tab, -10, -1, threads=THREADS)
    s32 = O.search32(q, codes, off, tab, -10, -1, threads=THREADS)
    fits = s32 <= 32767
    assert np.array_equal(s16[fits].astype(np.int32), s32[fits]), "reference != oracle where the score fits int16"
    print("c3: true scores", int(s32.min()), "..", int(s32.max()), "; > 32767:", int((~fits).sum()),
          "; beyond the engine's biased-u16 range:", int((s32 > 65535 - 129 - 140).sum()))
    np.savez_compressed(os.path.join(OUT, "c3_full.npz"), ref16=s16, true32=s32, n_residues=int(off[-1]))


def gen_c4():
    tab, _ = synth.load_matrix(os.path.join(ROOT, "scoring", "BLOSUM62.txt"))
    q, codes, off = synth.config4()
    idx = np.unique(np.concatenate([np.arange(64), np.arange(64, 1024, 17)])).astype(np.int64)
    c, o = take(codes, off, idx)
    s32 = O.search32(q, c, o, tab, -10, -1, threads=THREADS)
    print("c4:", len(idx), "targets, max", int(s32.max()), "second", int(np.sort(s32)[-2]))
    np.savez_compressed(os.path.join(OUT, "c4_sample.npz"), idx=idx, true32=s32, n_residues=int(off[-1]))


LETTERS = "?ABCDEFGHIJKLMNOPQRSTUVWXYZ????*"


def to_text(codes):
    return "".join(LETTERS[int(c)] for c in codes)


def planted_pair(g, k):
    """A query and a target sharing a mutated segment with a few indels; every fourth pair is unrelated background."""
    lq, lt = int(g.integers(20, 220)), int(g.integers(20, 260))
    q, t = synth.residues(g, lq), synth.residues(g, lt)
    if k % 4 != 3:
        m = int(g.integers(8, min(lq, lt) + 1))
        a, b = int(g.integers(0, lq - m + 1)), int(g.integers(0, lt - m + 1))
        seg = q[a:a + m].copy()
        mut = g.random(m) < g.uniform(0.0, 0.3)
        seg[mut] = synth.residues(g, int(mut.sum()))
        if m > 20 and k % 2 == 0:
            cut = int(g.integers(5, m - 5)); dl = int(g.integers(1, 4))
            seg = np.concatenate([seg[:cut], seg[cut + dl:]])
        t = np.concatenate([t[:b], seg, t[b + m:]])[:max(lt, 8)]
    return q, t


def gen_upstream():
    assert os.path.exists(UPSTREAM)
    settings = [("PAM250", None), ("BLOSUM62", (-10, -1))]
    rows = {k: [] for k in ("setting", "q", "t", "score", "q_pos", "q_len", "t_pos", "t_len", "q_aln", "t_aln")}
    g = synth.rng(2024)
    for si, (mat, gaps) in enumerate(settings):
        for k in range(104):
            q, t = planted_pair(g, k)
            cmd = [UPSTREAM, "--substitution_matrix", os.path.join(ROOT, "scoring", mat + ".txt"), "--maxhits", "1", "--minscore", "1"]
            if gaps:
                cmd += ["--gapopen", str(gaps[0]), "--gapextend", str(gaps[1])]
            out = subprocess.run(cmd + [to_text(q), to_text(t)], capture_output=True, text=True, check=True).stdout
            m = re.search(r"hit 0\.0 score: (-?\d+)\n\s+(\S+)\s+\[pos: (\d+); len: (\d+)\]\n\s+(\S+)\s+\[pos: (\d+); len: (\d+)\]", out)
            if m is None:                       # no local alignment above the minimum score
                vals = (0, -1, 0, -1, 0, "", "")
            else:
                vals = (int(m.group(1)), int(m.group(3)), int(m.group(4)), int(m.group(6)), int(m.group(7)), m.group(2), m.group(5))
            for key, v in zip(("score", "q_pos", "q_len", "t_pos", "t_len", "q_aln", "t_aln"), vals):
                rows[key].append(v)
            rows["setting"].append(si); rows["q"].append(to_text(q)); rows["t"].append(to_text(t))
    np.savez_compressed(os.path.join(OUT, "upstream_hits.npz"),
                        settings=np.array([f"{m}:{'default' if g is None else f'{g[0]},{g[1]}'}" for m, g in settings]),
                        **{k: np.array(v) for k, v in rows.items()})
    sc = np.array(rows["score"])
    print("upstream:", len(sc), "pairs; scores", int(sc.min()), "..", int(sc.max()), "; no hit:", int((sc == 0).sum()))


POS_GAPS = [(1, -3), (2, -1), (0, 1), (1, 1), (-5, 2), (3, 0)]


def gen_posgap():
    """Positive gap scores: the reference accepts them (src/alignment_cmdline.c:255-267) and runs its three-state recurrence
    on them (src/alignment.c:124-161).  With positive gap scores the reference's '*' padding rows are NOT neutral (E / F keep
    growing across them and feed M), so its result depends on the batch a target sits in -- the same class of quirk as
    SURVEY.md A.5.  Hence two sets:
      equal   17 batches of 16 EQUAL-length targets (no padding): reference == per-target truth; the engine must match;
      ragged  3 ordinary batches: the reference's padded computation, which swgpu_fill_batch16 / alignment_fill_matrices
              (the twins of that computation) must reproduce.
    Plus the reference CLI's stdout with --gapopen 1 --gapextend 1 on a database of two equal-length batches."""
    import tempfile
    tab, _ = synth.load_matrix(os.path.join(ROOT, "scoring", "BLOSUM62.txt"))
    g = synth.rng(606)
    q = synth.make_query(607, 300)

    def related(codes, off, step):
        for k in range(0, len(off) - 1, step):
            L = int(off[k + 1] - off[k]); src = np.resize(q, L).copy(); mut = g.random(L) < 0.2
            src[mut] = synth.residues(g, int(mut.sum())); codes[off[k]:off[k + 1]] = src

    lens = np.repeat(np.sort(g.integers(1, 180, size=17))[::-1], 16)
    off = np.zeros(len(lens) + 1, dtype=np.int64); np.cumsum(lens, out=off[1:])
    codes = synth.residues(g, off[-1]); related(codes, off, 4)
    rl = np.sort(g.integers(1, 150, size=48))[::-1]
    roff = np.zeros(49, dtype=np.int64); np.cumsum(rl, out=roff[1:])
    rcodes = synth.residues(g, roff[-1]); related(rcodes, roff, 3)
    res = {}
    for lq in (40, 130, 300):
        for go, ge in POS_GAPS:
            s16, _ = O.ref_search16(q[:lq], codes, off, tab, go, ge, threads=THREADS)
            s32 = O.search32(q[:lq], codes, off, tab, go, ge, threads=THREADS)
            assert int(s32.max()) < 32000 and np.array_equal(s16.astype(np.int32), s32), (lq, go, ge)
            res[f"s_{lq}_{go}_{ge}"] = s16
            r16, _ = O.ref_search16(q[:lq], rcodes, roff, tab, go, ge, threads=THREADS)
            pad = [O.score16_wrap(q[:lq], rcodes[roff[t]:roff[t + 1]], int(rl[t // 16 * 16]), tab, go, ge) for t in range(48)]
            assert list(r16) == pad, (lq, go, ge)               # the restatement of the padded arithmetic agrees
            res[f"r_{lq}_{go}_{ge}"] = r16
    print("posgap: max score", max(int(v.max()) for v in res.values()),
          "; ragged targets whose padded score differs from the unpadded one:",
          int(sum((res[f"r_300_{go}_{ge}"].astype(np.int32) != O.search32(q, rcodes, roff, tab, go, ge)).sum() for go, ge in POS_GAPS)))
    cl = np.repeat([70, 41], 16)
    coff = np.zeros(33, dtype=np.int64); np.cumsum(cl, out=coff[1:])
    ccodes = synth.residues(g, coff[-1]); related(ccodes, coff, 5)
    cq = q[:90]
    with tempfile.TemporaryDirectory() as td:
        qf, df = os.path.join(td, "q.fa"), os.path.join(td, "db.fa")
        synth.to_fasta(qf, cq, np.array([0, len(cq)]), prefix="query"); synth.to_fasta(df, ccodes, coff, prefix="sp")
        r = subprocess.run([O.REF_BIN, "--substitution_matrix", os.path.join(ROOT, "scoring", "BLOSUM62.txt"), "--gapopen", "1",
                            "--gapextend", "1", "--files", qf, df], capture_output=True, text=True, env={**os.environ, "OMP_NUM_THREADS": "2"})
        assert r.returncode == 0, r.stderr
        cli, qtxt, dtxt = r.stdout.replace(td, "TMP"), open(qf).read(), open(df).read()
    np.savez_compressed(os.path.join(OUT, "ref16_posgap.npz"), codes=codes, off=off, rcodes=rcodes, roff=roff, query=q, table=tab,
                        cli_stdout=cli, cli_query_fasta=qtxt, cli_db_fasta=dtxt, **res)


if __name__ == "__main__":
    assert O.have_ref(), "build oracle/_ref first (make -C oracle)"
    which = sys.argv[1:] or ["c2", "c3", "c4", "upstream", "posgap"]
    for w in which:
        {"c2": gen_c2, "c3": gen_c3, "c4": gen_c4, "upstream": gen_upstream, "posgap": gen_posgap}[w]()
