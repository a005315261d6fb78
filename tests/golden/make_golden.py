"""Generates tests/golden/*.npz from the UNMODIFIED reference compiled under oracle/_ref
(oracle/Makefile builds it from the sources where they lie in /root/reference).

Run in the dev container only (needs /root/reference):   python tests/golden/make_golden.py
The vectors hold inputs AND the reference's outputs, so the GPU box (no /root/reference) can replay them.

  ref16_<matrix>.npz   int16 scores of the reference's AVX2 kernel (alignment_fill_matrices via
                       oracle/ref_harness.c) for a seeded, length-sorted database, several queries and
                       gap settings.
  ref16_overflow.npz   PAM250 low-complexity targets whose true scores straddle 32 767: the reference's
                       wrapped int16 output beside the inputs.
  cli_small.npz        stdout of the reference's own bin/smith_waterman on a small FASTA pair
                       (grammar of src/tools/sw_cmdline.c:38-75) + the parsed scores.
  encode.npz           letters_to_index over all 256 byte values (-1 where the reference exit(1)s).
"""
import os, subprocess, sys, tempfile
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "seq-align-gpu_b200"))
from swgpu import synth            # noqa: E402
from oracle import oracle as O     # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
GAPS = [(-10, -1), (-2, -1), (0, -1), (-11, -1), (-5, -3), (-1, -1)]


def sorted_db(seed, n, lo, hi, median=120.0):
    g = synth.rng(seed)
    lens = np.sort(np.clip(np.floor(g.lognormal(np.log(median), 0.8, size=n)), lo, hi).astype(np.int64))[::-1].copy()
    off = np.zeros(n + 1, dtype=np.int64); np.cumsum(lens, out=off[1:])
    return synth.residues(g, off[-1]), off


def main():
    assert O.have_ref(), "build oracle/_ref first (make -C oracle)"
    # 1. per-matrix score vectors
    for mi, name in enumerate(("BLOSUM62", "BLOSUM45", "PAM250")):
        tab, diag = synth.load_matrix(os.path.join(ROOT, "scoring", name + ".txt"))
        codes, off = sorted_db(500 + mi, 272, 1, 900)        # 272 = 17 x 16 (the reference needs N % 16 == 0)
        # a few related targets (mutated copies of the query region) so scores are not all background
        q_all = synth.make_query(900 + mi, 700)
        g = synth.rng(77 + mi)
        for k in range(0, 272, 5):
            L = int(off[k + 1] - off[k])
            src = q_all[:L].copy() if L <= 700 else np.resize(q_all, L)
            mut = g.random(L) < 0.15
            src[mut] = synth.residues(g, int(mut.sum()))
            codes[off[k]:off[k + 1]] = src
        res = {}
        for lq in (1, 7, 33, 144, 360, 700):
            for (go, ge) in GAPS:
                s, _ = O.ref_search16(q_all[:lq], codes, off, tab, go, ge, threads=4)
                res[f"s_{lq}_{-go}_{-ge}"] = s
        np.savez_compressed(os.path.join(OUT, f"ref16_{name}.npz"), codes=codes, off=off, query=q_all, table=tab, **res)
        print(name, "max score", max(int(v.max()) for v in res.values()))

    # 2. overflow vectors (PAM250, W/C/Y/F low-complexity generator of SURVEY.md 8(d) C3)
    tab, _ = synth.load_matrix(os.path.join(ROOT, "scoring", "PAM250.txt"))
    g = synth.rng(44)
    lc_codes = np.array([23, 3, 25, 6], dtype=np.int8)       # W C Y F
    q = lc_codes[g.choice(4, size=3000, p=[0.4, 0.2, 0.2, 0.2])]
    lens = np.sort(g.integers(2300, 2701, size=48))[::-1]
    parts = []
    for L in lens:
        st = int(g.integers(0, 3000 - L + 1))
        parts.append(q[st:st + L])
    off = np.zeros(49, dtype=np.int64); np.cumsum(lens, out=off[1:])
    codes = np.concatenate(parts)
    s16, _ = O.ref_search16(q, codes, off, tab, -10, -1, threads=4)
    s32 = O.search32(q, codes, off, tab, -10, -1)
    np.savez_compressed(os.path.join(OUT, "ref16_overflow.npz"), codes=codes, off=off, query=q, table=tab, ref16=s16, true32=s32)
    print("overflow: true", int(s32.min()), "..", int(s32.max()), " ref16 exact for", int((s16 == s32).sum()), "of 48")

    # 3. the reference CLI on a small FASTA pair
    codes, off = sorted_db(321, 32, 5, 200, median=60.0)
    qq = synth.make_query(322, 90)
    with tempfile.TemporaryDirectory() as td:
        qf, df = os.path.join(td, "q.fa"), os.path.join(td, "db.fa")
        synth.to_fasta(qf, qq, np.array([0, len(qq)]), prefix="query")
        synth.to_fasta(df, codes, off, prefix="sp")
        outs = {}
        for tag, extra in (("plain", []), ("fasta", ["--printfasta"])):
            cmd = [O.REF_BIN, "--substitution_matrix", os.path.join(ROOT, "scoring", "BLOSUM62.txt"),
                   "--gapopen", "-10", "--gapextend", "-1", *extra, "--files", qf, df]
            r = subprocess.run(cmd, capture_output=True, text=True, env={**os.environ, "OMP_NUM_THREADS": "2"})
            assert r.returncode == 0, r.stderr
            outs[tag] = r.stdout.replace(td, "TMP")
        qtxt, dtxt = open(qf).read(), open(df).read()
    import re
    sc = np.array([int(m.group(2)) for m in re.finditer(r"Entry\s+#(\d+):\s*score:\s*([+-]?\d+)", outs["plain"])], dtype=np.int32)
    assert sc.size == 32
    np.savez_compressed(os.path.join(OUT, "cli_small.npz"), query_fasta=qtxt, db_fasta=dtxt, stdout_plain=outs["plain"],
                        stdout_fasta=outs["fasta"], scores=sc, codes=codes, off=off, query=qq)

    # 4. letters_to_index over every byte (run in a subprocess per illegal byte would exit(1); the reference
    #    function is pure for legal bytes, so legal ones are called directly and the rest recorded as -1)
    enc = np.full(256, -1, dtype=np.int32)
    for c in range(256):
        if chr(c).isascii() and (chr(c).isalpha() or chr(c) == "*"):
            enc[c] = O.ref().refh_letters_to_index(c)
    np.savez_compressed(os.path.join(OUT, "encode.npz"), enc=enc)
    print("golden vectors written to", OUT)


if __name__ == "__main__":
    main()
