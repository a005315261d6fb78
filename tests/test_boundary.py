"""The drop-in boundary, exercised the way a user of the reference would:

  * the reference's UNMODIFIED main, /root/reference/src/tools/sw_cmdline.c, compiles against this repo's include/ and
    links to libalign.so + libswgpu.so (oracle/Makefile builds it as oracle/_ref/ref_main_on_libalign in the dev
    container; the binary travels to the GPU box) -- and, on the GPU, prints what the reference's own binary printed;
  * aligner_create -> alignment_fill_matrices -> max_scores / max_scores32 -> aligner_update -> aligner_destroy through
    libalign.so, from several host threads at once: the reference's contract (alignment_fill_matrices is re-entrant
    across distinct aligner_t and called from every OpenMP worker, src/alignment_cmdline.c:504-507);
  * a reference-style driver loop over alignment_fill_matrices for the whole of C1 (625 batches) finishes in seconds.
"""
import ctypes as C
import os
import re
import subprocess
import threading
import time

import numpy as np
import pytest

from conftest import ROOT, SCORING, golden

REF = "/root/reference"
REF_MAIN = os.path.join(ROOT, "oracle", "_ref", "ref_main_on_libalign")
LIBALIGN = os.path.join(ROOT, "seq-align-gpu_b200", "libalign.so")


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "src")), reason="needs the reference tree (dev container only)")
def test_reference_main_compiles_and_links_against_this_repo():
    """gcc on the reference's own src/tools/sw_cmdline.c with -I include first (our alignment*.h win), its seq_file.h and
    alignment_macros.h from the reference tree, linked to libalign.so: no source change."""
    r = subprocess.run(["make", "-s", "-B", "-C", os.path.join(ROOT, "oracle"), "-f", os.path.join(ROOT, "oracle", "Makefile"),
                        REF_MAIN], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert os.path.exists(REF_MAIN)
    und = subprocess.run(["nm", "-D", "--undefined-only", REF_MAIN], capture_output=True, text=True).stdout
    for sym in ("cmdline_new", "align_from_query_and_db", "scoring_system_default", "cmdline_free"):
        assert sym in und, sym                                  # resolved by libalign.so at run time
    ldd = subprocess.run(["ldd", REF_MAIN], capture_output=True, text=True).stdout
    assert "libalign.so" in ldd and "libswgpu.so" in ldd and "not found" not in ldd
    # no GPU needed for the usage path: same text and exit status as the reference's
    u = subprocess.run([REF_MAIN], capture_output=True, text=True)
    assert u.returncode == 1 and u.stderr.startswith("usage:")


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["plain", "fasta"])
def test_reference_main_on_libalign_prints_what_the_reference_printed(tmp_path, mode):
    if not os.path.exists(REF_MAIN):
        pytest.skip("oracle/_ref/ref_main_on_libalign was not built (it is built in the dev container and travels)")
    g = golden("cli_small.npz")
    q = tmp_path / "q.fa"; q.write_text(str(g["query_fasta"]))
    d = tmp_path / "db.fa"; d.write_text(str(g["db_fasta"]))
    extra = ["--printfasta"] if mode == "fasta" else []
    r = subprocess.run([REF_MAIN, "--substitution_matrix", os.path.join(SCORING, "BLOSUM62.txt"), "--gapopen", "-10",
                        "--gapextend", "-1", *extra, "--files", str(q), str(d)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    norm = lambda s: re.sub(r"Total Time: [0-9.]+", "Total Time: T", s)
    assert norm(r.stdout.replace(str(tmp_path), "TMP")) == norm(str(g["stdout_" + mode]))


# ---- the aligner_t API through libalign.so (ctypes mirror of include/alignment.h / alignment_scoring.h) ----------------
class Scoring(C.Structure):
    _fields_ = [("gap_open", C.c_int16), ("gap_extend", C.c_int16), ("use_match_mismatch", C.c_bool),
                ("match", C.c_int), ("mismatch", C.c_int), ("case_sensitive", C.c_bool),
                ("swap_set", C.c_uint32 * 32), ("pad_to_32", C.c_uint8 * 12),       # alignas(32) swap_scores
                ("swap_scores", (C.c_int8 * 32) * 32), ("min_penalty", C.c_int), ("max_penalty", C.c_int),
                ("tail_pad", C.c_uint8 * 24)]


class Aligner(C.Structure):
    _fields_ = [("scoring", C.c_void_p), ("seq_a_indexes", C.c_void_p), ("seq_b_batch_indexes", C.c_void_p),
                ("seq_a_str", C.c_void_p), ("seq_b_str_batch", C.c_void_p), ("seq_a_fasta", C.c_void_p),
                ("seq_b_fasta_batch", C.c_void_p), ("vector_size", C.c_size_t), ("score_width", C.c_size_t),
                ("score_height", C.c_size_t), ("curr_match_scores", C.c_void_p), ("curr_gap_a_scores", C.c_void_p),
                ("curr_gap_b_scores", C.c_void_p), ("max_scores", C.POINTER(C.c_int16)), ("max_scores32", C.POINTER(C.c_int32))]


def libalign():
    L = C.CDLL(LIBALIGN)
    L.aligner_create.restype = C.POINTER(Aligner)
    L.aligner_create.argtypes = [C.c_void_p] * 6 + [C.c_size_t] * 3 + [C.c_void_p]
    L.aligner_update.restype = None
    L.aligner_update.argtypes = [C.POINTER(Aligner)] + [C.c_void_p] * 6 + [C.c_size_t] * 3 + [C.c_void_p]
    L.alignment_fill_matrices.restype = None
    L.alignment_fill_matrices.argtypes = [C.POINTER(Aligner)]
    L.aligner_destroy.restype = None
    L.aligner_destroy.argtypes = [C.POINTER(Aligner)]
    L.scoring_init.restype = None
    L.scoring_init.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_bool]
    return L


def test_ctypes_mirrors_have_the_reference_layout():
    """the ctypes structs above against the layout pinned in tests/test_host_layer.py (1216 / 160 / 104)"""
    assert C.sizeof(Scoring) == 1216 and Scoring.swap_scores.offset == 160
    assert Aligner.max_scores.offset == 104 and Aligner.max_scores32.offset == 112


def make_batches(codes, off, n_batches):
    """the reference's batch build (src/alignment_cmdline.c:429-459): 16 consecutive sorted targets, [pos][lane], '*' = 31 padding"""
    out = []
    for b in range(n_batches):
        first = b * 16
        lanes = min(16, len(off) - 1 - first)
        plen = int(off[first + 1] - off[first])
        buf = np.full((max(plen, 1), 16), 31, dtype=np.int8)
        for l in range(lanes):
            t = codes[off[first + l]:off[first + l + 1]]
            buf[:len(t), l] = t
        out.append((np.ascontiguousarray(buf.reshape(-1)), plen, lanes))
    return out


@pytest.mark.gpu
def test_aligner_api_through_libalign_from_four_threads(oracle, matrices):
    from swgpu import synth
    L = libalign()
    tab = np.ascontiguousarray(matrices["BLOSUM62"], dtype=np.int8)
    sc = Scoring()
    L.scoring_init(C.byref(sc), 1, -1, -10, -1, True)
    C.memmove(C.byref(sc, Scoring.swap_scores.offset), tab.ctypes.data, 1024)
    q, codes, off = synth.config1()
    nb = 48
    batches = make_batches(codes, off, nb)
    want = [[oracle.score16_wrap(q, codes[off[b * 16 + l]:off[b * 16 + l + 1]], plen, tab, -10, -1) for l in range(lanes)]
            for b, (_, plen, lanes) in enumerate(batches)]
    got = [None] * nb
    errors = []

    def worker(tid):
        try:
            a = None
            for b in range(tid, nb, 4):
                buf, plen, lanes = batches[b]
                if a is None:                                 # first batch of this thread creates, later ones re-point (as the reference's driver does)
                    a = L.aligner_create(None, None, None, None, q.ctypes.data, buf.ctypes.data, len(q), plen, lanes, C.addressof(sc))
                else:
                    L.aligner_update(a, None, None, None, None, q.ctypes.data, buf.ctypes.data, len(q), plen, lanes, C.addressof(sc))
                L.alignment_fill_matrices(a)
                got[b] = ([int(a.contents.max_scores[l]) for l in range(lanes)], [int(a.contents.max_scores32[l]) for l in range(lanes)])
            if a is not None:
                L.aligner_destroy(a)
                # like the reference, destroy leaves max_scores to the caller: still readable
                assert int(a.contents.max_scores[0]) == got[b][0][0]
        except Exception as e:                                # noqa: BLE001
            errors.append(repr(e))

    th = [threading.Thread(target=worker, args=(t,)) for t in range(4)]
    [t.start() for t in th]; [t.join() for t in th]
    assert not errors, errors
    for b in range(nb):
        assert got[b][0] == want[b] and got[b][1] == want[b], (b, got[b], want[b])


@pytest.mark.gpu
def test_reference_style_driver_loop_over_c1_finishes_in_seconds(oracle, matrices):
    """A user who keeps the reference's driver loop: all 625 batches of C1 through alignment_fill_matrices, one thread.
    (Round 1 built and destroyed a whole engine handle per call; now each host thread keeps one context.)"""
    from swgpu import synth
    L = libalign()
    tab = np.ascontiguousarray(matrices["BLOSUM62"], dtype=np.int8)
    sc = Scoring()
    L.scoring_init(C.byref(sc), 1, -1, -10, -1, True)
    C.memmove(C.byref(sc, Scoring.swap_scores.offset), tab.ctypes.data, 1024)
    q, codes, off = synth.config1()
    batches = make_batches(codes, off, 625)
    a = L.aligner_create(None, None, None, None, q.ctypes.data, batches[0][0].ctypes.data, len(q), batches[0][1], 16, C.addressof(sc))
    L.alignment_fill_matrices(a)                              # first call: CUDA context + buffers
    t0 = time.perf_counter()
    got = np.zeros(10000, dtype=np.int32)
    for b, (buf, plen, lanes) in enumerate(batches):
        L.aligner_update(a, None, None, None, None, q.ctypes.data, buf.ctypes.data, len(q), plen, lanes, C.addressof(sc))
        L.alignment_fill_matrices(a)
        got[b * 16:b * 16 + lanes] = [a.contents.max_scores32[l] for l in range(lanes)]
    secs = time.perf_counter() - t0
    L.aligner_destroy(a)
    want = oracle.search32(q, codes, off, tab, -10, -1)       # no '*' in the query: padding rows cannot change a score
    assert np.array_equal(got, want)
    print(f"625 alignment_fill_matrices calls (C1, one host thread): {secs:.3f} s = {secs / 625 * 1e6:.0f} us per batch")
    assert secs < 10.0
