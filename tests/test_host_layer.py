"""Host C layer (libalign.so + bin/smith_waterman): the reference's own API names and CLI grammar.
CPU tests cover parsing, encoding, flags and error behaviour; the GPU test replays the reference CLI's
recorded stdout (tests/golden/cli_small.npz, produced by the unmodified reference binary)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT, SCORING, golden

BIN = os.environ.get("SWGPU_HOST_BIN") or os.path.join(ROOT, "bin", "smith_waterman")
HOST = os.environ.get("SWGPU_HOST_LIB") or os.path.join(ROOT, "seq-align-gpu_b200", "libalign.so")   # scripts/host_asan.sh: the ASan build


class Scoring(C.Structure):          # include/alignment_scoring.h (== the reference's scoring_t layout)
    _fields_ = [("gap_open", C.c_int16), ("gap_extend", C.c_int16), ("use_match_mismatch", C.c_bool),
                ("match", C.c_int), ("mismatch", C.c_int), ("case_sensitive", C.c_bool),
                ("swap_set", C.c_uint32 * 32), ("_pad", C.c_char * 12), ("swap_scores", (C.c_int8 * 32) * 32),
                ("min_penalty", C.c_int), ("max_penalty", C.c_int)]


def host():
    L = C.CDLL(HOST)
    L.letters_to_index.restype = C.c_int8
    L.letters_to_index.argtypes = [C.c_char]
    L.index_to_letters.restype = C.c_char
    L.index_to_letters.argtypes = [C.c_int8]
    return L


def load_with_host(path):
    L = host()
    z = C.CDLL("libz.so.1")
    z.gzopen.restype = C.c_void_p
    z.gzopen.argtypes = [C.c_char_p, C.c_char_p]
    z.gzclose.argtypes = [C.c_void_p]
    buf = C.create_string_buffer(C.sizeof(Scoring) + 64)
    addr = (C.addressof(buf) + 31) // 32 * 32            # alignas(32)
    sc = Scoring.from_address(addr)
    L.scoring_init(C.byref(sc), 1, -1, -10, -1, True)
    f = z.gzopen(path.encode(), b"r")
    assert f
    L.align_scoring_load_matrix.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_char]
    L.align_scoring_load_matrix(f, path.encode(), C.byref(sc), b"\x01")
    z.gzclose(f)
    return sc, buf


def test_struct_layouts_match_reference(tmp_path):
    """sizeof/offsetof measured on the reference's own headers (gcc, x86-64): scoring_t 1216 bytes, swap_set @20,
    swap_scores @160 (alignas 32), min_penalty @1184; aligner_t: max_scores @104 (ours appends max_scores32)."""
    assert Scoring.swap_scores.offset == 160 and Scoring.swap_set.offset == 20 and Scoring.min_penalty.offset == 1184
    src = tmp_path / "lay.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "alignment_cmdline.h"\n'
                   'int main(){printf("%zu %zu %zu %zu %zu %zu %zu", sizeof(scoring_t), offsetof(scoring_t,swap_set),'
                   'offsetof(scoring_t,swap_scores), offsetof(scoring_t,min_penalty), offsetof(aligner_t,vector_size),'
                   'offsetof(aligner_t,max_scores), offsetof(aligner_t,max_scores32));return 0;}\n')
    exe = tmp_path / "lay"
    subprocess.run(["/usr/bin/gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True).stdout.split()
    assert [int(x) for x in out] == [1216, 20, 160, 1184, 56, 104, 112]


@pytest.mark.parametrize("name", ["BLOSUM62", "BLOSUM45", "PAM250"])
def test_matrix_loader_equals_reference_tables(name):
    sc, _keep = load_with_host(os.path.join(SCORING, name + ".txt"))
    tab = np.array([[sc.swap_scores[a][b] for b in range(32)] for a in range(32)], dtype=np.int8)
    assert np.array_equal(tab, golden(f"ref16_{name}.npz")["table"])
    assert sc.gap_open == -10 and sc.gap_extend == -1
    # diagonal swap bits: defined for the 24 matrix letters only (J, O, U are not)
    defined = [bool((sc.swap_set[i] >> i) & 1) for i in range(32)]
    assert defined[ord("A") - 64] and defined[31] and not defined[ord("J") - 64] and not defined[ord("U") - 64]


def test_matrix_loader_separator_format_and_gzip(tmp_path):
    import gzip
    p = tmp_path / "m.txt.gz"
    with gzip.open(p, "wt") as f:
        f.write("# comment\n,A,C,G\nA,5,-1,-2\n\nC,-1,6,-3\nG,-2,-3,7\n")
    sc, _keep = load_with_host(str(p))
    assert sc.swap_scores[1][1] == 5 and sc.swap_scores[3][7] == -3 and sc.swap_scores[7][7] == 7
    assert sc.swap_scores[2][2] == 0                       # zero-initialised table (documented deviation)


def test_letters_to_index_matches_reference_golden():
    L = host()
    enc = golden("encode.npz")["enc"]
    for c in range(256):
        if enc[c] >= 0:
            assert L.letters_to_index(bytes([c])) == enc[c]
    assert L.index_to_letters(1) == b"A" and L.index_to_letters(26) == b"Z" and L.index_to_letters(31) == b"*"


def run(args, **kw):
    return subprocess.run([BIN, *args], capture_output=True, text=True, **kw)


def test_cli_usage_errors_like_reference():
    r = run([])
    assert r.returncode == 1 and r.stderr.startswith("usage:")
    r = run(["--files", "only_one"])
    assert r.returncode == 1 and "Error:" in r.stderr
    r = run(["--gapopen", "abc", "--files", "a", "b"])
    assert r.returncode == 1 and "Invalid --gapopen argument ('abc') must be an int" in r.stderr
    r = run(["--substitution_matrix", "/nonexistent", "--files", "a", "b"])
    assert r.returncode == 1 and "Couldn't read: /nonexistent" in r.stderr
    r = run(["--nonsense", "x", "--files", "a", "b"])
    assert r.returncode == 1 and "Unknown argument '--nonsense'" in r.stderr


def test_cli_file_errors_and_illegal_residue(tmp_path):
    m = os.path.join(SCORING, "BLOSUM62.txt")
    db = tmp_path / "db.fa"; db.write_text(">t\nACDEF\n")
    r = run(["--substitution_matrix", m, "--files", str(tmp_path / "missing.fa"), str(db)])
    assert r.returncode == 0 and "Error: couldn't open query file" in r.stderr           # exit status stays 0, like the reference
    assert r.stdout.startswith(f"Query File={tmp_path / 'missing.fa'} and Database File={db}\n")
    empty = tmp_path / "e.fa"; empty.write_text("")
    r = run(["--substitution_matrix", m, "--files", str(empty), str(db)])
    assert r.returncode == 0 and "is empty or invalid" in r.stderr
    bad = tmp_path / "bad.fa"; bad.write_text(">q\nAC-DE\n")
    r = run(["--substitution_matrix", m, "--files", str(bad), str(db)])
    assert r.returncode == 1 and "Error: - is not a legal character for the substitution matrix!" in r.stdout


def _write_cli_inputs(tmp_path):
    g = golden("cli_small.npz")
    q = tmp_path / "q.fa"; q.write_text(str(g["query_fasta"]))
    d = tmp_path / "db.fa"; d.write_text(str(g["db_fasta"]))
    return g, q, d


def _normalise(out, tmp):
    out = out.replace(str(tmp), "TMP")
    return re.sub(r"Total Time: [0-9.]+", "Total Time: T", out)


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["plain", "fasta"])
def test_cli_stdout_equals_reference_binary(tmp_path, mode):
    g, q, d = _write_cli_inputs(tmp_path)
    extra = ["--printfasta"] if mode == "fasta" else []
    r = run(["--substitution_matrix", os.path.join(SCORING, "BLOSUM62.txt"), "--gapopen", "-10", "--gapextend", "-1",
             *extra, "--files", str(q), str(d)])
    assert r.returncode == 0, r.stderr
    want = re.sub(r"Total Time: [0-9.]+", "Total Time: T", str(g["stdout_" + mode]))
    assert _normalise(r.stdout, tmp_path) == want


@pytest.mark.gpu
def test_cli_unsorted_db_partial_batch_gzip_and_topk(tmp_path, oracle, matrices):
    """What the reference cannot do: unsorted database, N % 16 != 0, gz input; plus the top-k mode."""
    import gzip
    from swgpu import synth
    g = synth.rng(5)
    lens = g.integers(1, 400, size=37)
    off = np.zeros(38, np.int64); np.cumsum(lens, out=off[1:])
    codes = synth.residues(g, off[-1])
    qq = synth.make_query(6, 120)
    codes[off[20]:off[20] + 100] = qq[10:110]
    synth.to_fasta(str(tmp_path / "db.fa"), codes, off)
    with open(tmp_path / "db.fa", "rb") as f, gzip.open(tmp_path / "db.fa.gz", "wb") as z:
        z.write(f.read())
    synth.to_fasta(str(tmp_path / "q.fa"), qq, np.array([0, 120]), prefix="query")
    args = ["--substitution_matrix", os.path.join(SCORING, "PAM250.txt"), "--gapopen", "-10", "--gapextend", "-1",
            "--files", str(tmp_path / "q.fa"), str(tmp_path / "db.fa.gz")]
    r = run(args)
    assert r.returncode == 0, r.stderr
    got = [int(m.group(2)) for m in re.finditer(r"Entry\s+#(\d+):\s*score:\s*([+-]?\d+)", r.stdout)]
    want = oracle.search32(qq, codes, off, matrices["PAM250"], -10, -1)
    assert got == list(want) and "Total Entries: 37" in r.stdout
    r = run(args, env={**os.environ, "SWGPU_TOPK": "3"})
    hits = re.findall(r"Hit #(\d+): entry (\d+) score (\d+) query \[(\d+)\.\.(\d+)\] target \[(\d+)\.\.(\d+)\]", r.stdout)
    assert len(hits) == 3 and int(hits[0][1]) == 20 and int(hits[0][2]) == int(want.max())
    s, c = oracle.end_start(qq, codes[off[20]:off[21]], matrices["PAM250"], -10, -1)
    assert tuple(int(x) for x in hits[0][3:]) == c
    # SWGPU_CIGAR=1 adds the alignment path of every hit
    r = run(args, env={**os.environ, "SWGPU_TOPK": "3", "SWGPU_CIGAR": "1"})
    cigs = re.findall(r"  cigar: (\S*)", r.stdout)
    t20 = codes[off[20]:off[21]]
    want_cig, sc = oracle.traceback(qq[c[0]:c[1] + 1], t20[c[2]:c[3] + 1], matrices["PAM250"], -10, -1)
    assert len(cigs) == 3 and cigs[0] == want_cig and sc == s


@pytest.mark.gpu
def test_cli_many_queries(tmp_path, oracle, matrices):
    """Many-to-many front end: SWGPU_ALL_QUERIES=1 aligns every record of the query file; by
    default only the first one is used, like the reference (src/alignment_cmdline.c:369-380)."""
    from swgpu import synth
    g = synth.rng(15)
    lens = g.integers(1, 300, size=53)
    off = np.zeros(54, np.int64); np.cumsum(lens, out=off[1:])
    codes = synth.residues(g, off[-1])
    qlens = [80, 1, 300, 33, 150]
    qoff = np.zeros(len(qlens) + 1, np.int64); np.cumsum(qlens, out=qoff[1:])
    qcodes = synth.residues(g, qoff[-1])
    codes[off[7]:off[7] + 60] = qcodes[qoff[2] + 20: qoff[2] + 80]
    synth.to_fasta(str(tmp_path / "db.fa"), codes, off)
    synth.to_fasta(str(tmp_path / "q.fa"), qcodes, qoff, prefix="query")
    tab = matrices["BLOSUM62"]
    args = ["--substitution_matrix", os.path.join(SCORING, "BLOSUM62.txt"), "--gapopen", "-10", "--gapextend", "-1",
            "--files", str(tmp_path / "q.fa"), str(tmp_path / "db.fa")]
    pat = r"Entry\s+#(\d+):\s*score:\s*([+-]?\d+)"
    r = run(args)
    assert r.returncode == 0, r.stderr
    got = [int(m.group(2)) for m in re.finditer(pat, r.stdout)]
    assert got == list(oracle.search32(qcodes[:80], codes, off, tab, -10, -1))
    r = run(args, env={**os.environ, "SWGPU_ALL_QUERIES": "1"})
    assert r.returncode == 0, r.stderr
    got = [int(m.group(2)) for m in re.finditer(pat, r.stdout)]
    want = []
    for k in range(len(qlens)):
        want += list(oracle.search32(qcodes[qoff[k]:qoff[k + 1]], codes, off, tab, -10, -1))
    assert got == want and r.stdout.count("Total Entries: 53") == 1
    r = run(args, env={**os.environ, "SWGPU_ALL_QUERIES": "1", "SWGPU_TOPK": "2"})
    assert r.returncode == 0, r.stderr
    blocks = r.stdout.split("Query #")[1:]
    assert len(blocks) == len(qlens)
    top = re.findall(r"Hit #0: entry (\d+) score (\d+)", blocks[2])
    s2 = oracle.search32(qcodes[qoff[2]:qoff[3]], codes, off, tab, -10, -1)
    assert top and int(top[0][0]) == 7 and int(top[0][1]) == int(s2.max())


@pytest.mark.gpu
def test_cli_packed_db_cache(tmp_path):
    """SWGPU_DB_CACHE=1 writes <db>.swdb on the first run and mmaps it on the next; the listing
    (with --printseq, so names and sequences come from the cache too) is byte-identical."""
    g, q, d = _write_cli_inputs(tmp_path)
    args = ["--substitution_matrix", os.path.join(SCORING, "BLOSUM62.txt"), "--gapopen", "-10", "--gapextend", "-1",
            "--printfasta", "--printseq", "--files", str(q), str(d)]
    env = {**os.environ, "SWGPU_DB_CACHE": "1", "SWGPU_VERBOSE": "1"}
    r0 = run(args)
    r1 = run(args, env=env)
    assert r1.returncode == 0 and os.path.exists(str(d) + ".swdb") and "packed cache" not in r1.stderr
    r2 = run(args, env=env)
    assert r2.returncode == 0 and "packed cache, mmap" in r2.stderr
    assert _normalise(r0.stdout, tmp_path) == _normalise(r1.stdout, tmp_path) == _normalise(r2.stdout, tmp_path)
    # touching the source invalidates the cache
    os.utime(d, ns=(0, 10**18))
    r3 = run(args, env=env)
    assert r3.returncode == 0 and "packed cache" not in r3.stderr
    assert _normalise(r3.stdout, tmp_path) == _normalise(r0.stdout, tmp_path)
