"""Host database loader (host/db_loader.c): the multi-threaded mmap FASTA parser must return exactly what the
streaming reader returns (record semantics of SURVEY.md A.4), for any thread count; the packed .swdb cache must
round-trip and refuse stale or foreign files.  CPU only (no GPU call is made)."""
import ctypes as C
import gzip
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, SCORING

HOST = os.environ.get("SWGPU_HOST_LIB") or os.path.join(ROOT, "seq-align-gpu_b200", "libalign.so")   # scripts/host_asan.sh: the ASan build
BIN = os.environ.get("SWGPU_HOST_BIN") or os.path.join(ROOT, "bin", "smith_waterman")


class Swdb(C.Structure):
    _fields_ = [("n", C.c_int64), ("n_residues", C.c_int64), ("codes", C.POINTER(C.c_int8)),
                ("off", C.POINTER(C.c_int64)), ("names", C.POINTER(C.c_char)), ("name_off", C.POINTER(C.c_int64)),
                ("text", C.POINTER(C.c_char)), ("from_cache", C.c_int), ("threads_used", C.c_int),
                ("map", C.c_void_p), ("map_len", C.c_size_t)]


def lib():
    L = C.CDLL(HOST)
    L.swdb_load.restype = C.c_int
    L.swdb_load.argtypes = [C.c_char_p, C.c_bool, C.c_int, C.POINTER(Swdb)]
    L.swdb_cache_write.restype = C.c_int
    L.swdb_cache_write.argtypes = [C.c_char_p, C.POINTER(Swdb), C.c_char_p]
    L.swdb_cache_open.restype = C.c_int
    L.swdb_cache_open.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(Swdb)]
    L.swdb_free.argtypes = [C.POINTER(Swdb)]
    return L


def snapshot(d):
    n, r = d.n, d.n_residues
    off = np.ctypeslib.as_array(d.off, shape=(n + 1,)).copy()
    noff = np.ctypeslib.as_array(d.name_off, shape=(n + 1,)).copy()
    codes = np.ctypeslib.as_array(d.codes, shape=(max(r, 1),))[:r].copy()
    names = C.string_at(d.names, int(noff[-1])) if noff[-1] else b""
    text = C.string_at(d.text, r) if d.text and r else None
    return dict(n=n, off=off, noff=noff, codes=codes, names=names, text=text)


def make_fasta(seed, n_records, crlf_every=7, width_choices=(60, 61, 80, 1000)):
    g = np.random.default_rng(seed)
    letters = np.frombuffer(b"ACDEFGHIKLMNPQRSTVWYacdxBZX*", dtype=np.uint8)
    out = [b"\n\r\n"]                                   # blank lines before the first record
    want_names, want_seqs = [], []
    for i in range(n_records):
        L = int(g.integers(0, 2500)) if i % 50 else 0   # some empty records
        seq = letters[g.integers(0, len(letters), size=L)].tobytes()
        name = b"sp|%06d|X_%d  some description > with a bracket" % (i, L)
        eol = b"\r\n" if i % crlf_every == 0 else b"\n"
        out.append(b">" + name + eol)
        w = int(width_choices[i % len(width_choices)])
        for p in range(0, L, w):
            out.append(seq[p:p + w] + eol)
        if i % 11 == 0:
            out.append(eol)                             # blank line inside / after a record
        want_names.append(name); want_seqs.append(seq)
    data = b"".join(out)
    return data[:-1] if data.endswith(b"\n") else data, want_names, want_seqs   # no final newline


def test_parallel_parser_equals_streaming_reader(tmp_path):
    L = lib()
    data, names, seqs = make_fasta(3, 4000)
    assert len(data) > (1 << 20)                        # big enough for the multi-threaded path
    p = tmp_path / "db.fa"; p.write_bytes(data)
    pz = tmp_path / "db.fa.gz"
    with gzip.open(pz, "wb") as z:
        z.write(data)
    snaps = {}
    for key, path, thr in (("t1", p, 1), ("t3", p, 3), ("t8", p, 8), ("t32", p, 32), ("gz", pz, 4)):
        d = Swdb()
        assert L.swdb_load(str(path).encode(), True, thr, C.byref(d)) == 0
        snaps[key] = snapshot(d)
        if key == "t8":
            assert d.threads_used > 1
        if key == "gz":
            assert d.threads_used == 0                  # streaming reader
        L.swdb_free(C.byref(d))
    ref = snaps["gz"]
    assert ref["n"] == len(names)
    assert ref["names"] == b"".join(n + b"\0" for n in names)
    assert ref["text"] == b"".join(seqs)
    lut = np.full(256, -1, np.int8)
    for c in range(26):
        lut[ord("a") + c] = lut[ord("A") + c] = c + 1
    lut[ord("*")] = 31
    assert np.array_equal(ref["codes"], lut[np.frombuffer(ref["text"], np.uint8)])
    for key in ("t1", "t3", "t8", "t32"):
        s = snaps[key]
        assert s["n"] == ref["n"] and s["names"] == ref["names"] and s["text"] == ref["text"], key
        assert np.array_equal(s["off"], ref["off"]) and np.array_equal(s["noff"], ref["noff"]), key
        assert np.array_equal(s["codes"], ref["codes"]), key


def test_small_and_degenerate_files(tmp_path):
    L = lib()
    cases = {"empty.fa": b"", "blank.fa": b"\n\n", "one.fa": b">a\nACD", "hdr_only.fa": b">only header\n",
             "plain.txt": b"ACDE\nFGH\n"}
    want = {"empty.fa": (0, 0), "blank.fa": (0, 0), "one.fa": (1, 3), "hdr_only.fa": (1, 0), "plain.txt": (2, 7)}
    for name, data in cases.items():
        p = tmp_path / name; p.write_bytes(data)
        d = Swdb()
        assert L.swdb_load(str(p).encode(), True, 4, C.byref(d)) == 0, name
        assert (d.n, d.n_residues) == want[name], name
        L.swdb_free(C.byref(d))
    d = Swdb()
    assert L.swdb_load(str(tmp_path / "missing.fa").encode(), True, 4, C.byref(d)) == -1


def test_cache_roundtrip_and_staleness(tmp_path):
    L = lib()
    data, names, seqs = make_fasta(5, 300)
    p = tmp_path / "db.fa"; p.write_bytes(data)
    d = Swdb()
    assert L.swdb_load(str(p).encode(), True, 2, C.byref(d)) == 0
    a = snapshot(d)
    cache = str(tmp_path / "db.fa.swdb").encode()
    assert L.swdb_cache_write(cache, C.byref(d), str(p).encode()) == 0
    L.swdb_free(C.byref(d))
    c = Swdb()
    assert L.swdb_cache_open(cache, str(p).encode(), C.byref(c)) == 0
    b = snapshot(c)
    assert c.from_cache == 1 and b["text"] is None
    assert b["n"] == a["n"] and b["names"] == a["names"]
    assert np.array_equal(a["off"], b["off"]) and np.array_equal(a["codes"], b["codes"]) and np.array_equal(a["noff"], b["noff"])
    L.swdb_free(C.byref(c))
    # the source changed -> the cache is stale
    st = os.stat(p)
    with open(p, "ab") as f:
        f.write(b"\n>extra\nAAA\n")
    assert L.swdb_cache_open(cache, str(p).encode(), C.byref(c)) == -1
    # same size, different mtime -> stale as well
    p.write_bytes(data); os.utime(p, ns=(st.st_atime_ns, st.st_mtime_ns + 5_000_000_000))
    assert L.swdb_cache_open(cache, str(p).encode(), C.byref(c)) == -1
    os.utime(p, ns=(st.st_atime_ns, st.st_mtime_ns))
    assert L.swdb_cache_open(cache, str(p).encode(), C.byref(c)) == 0
    L.swdb_free(C.byref(c))
    # truncated / foreign files are refused
    blob = open(cache, "rb").read()
    (tmp_path / "trunc.swdb").write_bytes(blob[:-64])
    (tmp_path / "magic.swdb").write_bytes(b"XXXX" + blob[4:])
    assert L.swdb_cache_open(str(tmp_path / "trunc.swdb").encode(), None, C.byref(c)) == -1
    assert L.swdb_cache_open(str(tmp_path / "magic.swdb").encode(), None, C.byref(c)) == -1
    assert L.swdb_cache_open(str(tmp_path / "nope.swdb").encode(), None, C.byref(c)) == -1


def test_illegal_residue_reports_first_in_file_order(tmp_path):
    """Two illegal characters in different thread ranges: the message names the one that comes first in the file,
    on stdout with exit status 1 like the reference (src/alignment_scoring.c:78-79)."""
    data, _, _ = make_fasta(9, 3000)
    data = bytearray(data)
    cut = data.index(b"\n>", len(data) // 4)
    seq_a = data.index(b"\n", cut + 1) + 1              # first sequence byte of a record in the 2nd quarter
    while data[seq_a] in b">\r\n":
        seq_a = data.index(b"\n", seq_a + 1) + 1
    cut = data.index(b"\n>", 3 * len(data) // 4)
    seq_b = data.index(b"\n", cut + 1) + 1
    while data[seq_b] in b">\r\n":
        seq_b = data.index(b"\n", seq_b + 1) + 1
    data[seq_a] = ord("#"); data[seq_b] = ord("-")
    p = tmp_path / "bad.fa"; p.write_bytes(bytes(data))
    q = tmp_path / "q.fa"; q.write_text(">q\nACDEFGHIK\n")
    r = subprocess.run([BIN, "--substitution_matrix", os.path.join(SCORING, "BLOSUM62.txt"), "--files", str(q), str(p)],
                       capture_output=True, text=True, env={**os.environ, "SWGPU_PARSE_THREADS": "8"})
    assert r.returncode == 1
    assert "Error: # is not a legal character for the substitution matrix!" in r.stdout
