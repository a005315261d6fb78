"""ctypes access to the CPU checkers (TEST INFRASTRUCTURE: import only from tests/, smoke(), bench cpu_baseline).

  liboracle.so          -- the C restatement (oracle/sw_oracle.c)
  _ref/libalign_ref.so  -- the unmodified reference sources + oracle/ref_harness.c (built in the dev container)
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libalign_ref.so")
REF_BIN = os.path.join(HERE, "_ref", "smith_waterman")


def build():
    subprocess.run(["make", "-s", "-C", HERE], check=True)


_o = None
_r = None


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def oracle():
    global _o
    if _o is None:
        if not os.path.exists(ORACLE_SO):
            build()
        L = C.CDLL(ORACLE_SO)
        L.swo_score32.restype = C.c_int32
        L.swo_score32.argtypes = [C.c_void_p, C.c_long, C.c_void_p, C.c_long, C.c_void_p, C.c_int, C.c_int]
        L.swo_score16_wrap.restype = C.c_int16
        L.swo_score16_wrap.argtypes = [C.c_void_p, C.c_long, C.c_void_p, C.c_long, C.c_long, C.c_void_p, C.c_int, C.c_int]
        L.swo_end_start.restype = C.c_int32
        L.swo_end_start.argtypes = [C.c_void_p, C.c_long, C.c_void_p, C.c_long, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.swo_search32.restype = None
        L.swo_search32.argtypes = [C.c_void_p, C.c_long, C.c_void_p, C.c_void_p, C.c_long, C.c_void_p, C.c_int, C.c_int,
                                   C.c_int, C.c_void_p]
        L.swo_letters_to_index.restype = C.c_int
        L.swo_letters_to_index.argtypes = [C.c_int]
        L.swo_max_threads.restype = C.c_int
        _o = L
    return _o


def have_ref():
    return os.path.exists(REF_SO)


def ref():
    global _r
    if _r is None:
        L = C.CDLL(REF_SO)
        L.refh_search16.restype = C.c_double
        L.refh_search16.argtypes = [C.c_void_p, C.c_long, C.c_void_p, C.c_void_p, C.c_long, C.c_void_p, C.c_int, C.c_int,
                                    C.c_int, C.c_void_p]
        L.refh_max_threads.restype = C.c_int
        L.refh_letters_to_index.restype = C.c_int
        L.refh_letters_to_index.argtypes = [C.c_int]
        _r = L
    return _r


def _prep(q, codes, off, table):
    q = np.ascontiguousarray(q, dtype=np.int8)
    codes = np.ascontiguousarray(codes, dtype=np.int8)
    off = np.ascontiguousarray(off, dtype=np.int64)
    table = np.ascontiguousarray(np.asarray(table).reshape(-1), dtype=np.int8)
    return q, codes, off, table


def score32(q, t, table, gap_open, gap_extend):
    q, t, _, table = _prep(q, t, [0], table)
    return int(oracle().swo_score32(_p(q), q.size, _p(t), t.size, _p(table), gap_open, gap_extend))


def score16_wrap(q, t, padded_len, table, gap_open, gap_extend):
    q, t, _, table = _prep(q, t, [0], table)
    return int(oracle().swo_score16_wrap(_p(q), q.size, _p(t), t.size, padded_len, _p(table), gap_open, gap_extend))


def end_start(q, t, table, gap_open, gap_extend):
    q, t, _, table = _prep(q, t, [0], table)
    c = np.zeros(4, dtype=np.int32)
    s = int(oracle().swo_end_start(_p(q), q.size, _p(t), t.size, _p(table), gap_open, gap_extend, _p(c)))
    return s, tuple(int(x) for x in c)


def count_best(q, t, table, gap_open, gap_extend):
    """Number of cells attaining the best score (1 = unique end cell)."""
    q, t, _, table = _prep(q, t, [0], table)
    L = oracle()
    L.swo_count_best.restype = C.c_long
    return int(L.swo_count_best(_p(q), C.c_long(q.size), _p(t), C.c_long(t.size), _p(table), C.c_int(gap_open), C.c_int(gap_extend)))


def traceback(q, t, table, gap_open, gap_extend):
    """CIGAR of the forced-corner global alignment of q and t (the sub-sequences of a hit). Returns (cigar, score)."""
    q, t, _, table = _prep(q, t, [0], table)
    L = oracle()
    L.swo_traceback.restype = C.c_long
    ops = np.zeros(q.size + t.size + 1, dtype=np.uint32)
    sc = C.c_int32(0)
    n = L.swo_traceback(_p(q), C.c_long(q.size), _p(t), C.c_long(t.size), _p(table), C.c_int(gap_open), C.c_int(gap_extend),
                        _p(ops), C.c_long(ops.size), C.byref(sc))
    if n < 0:
        raise RuntimeError("oracle traceback failed")
    return "".join(f"{int(x) >> 2}{'MID'[int(x) & 3]}" for x in ops[:n]), int(sc.value)


def cigar_score(cigar, q, t, table, gap_open, gap_extend):
    """Score of a CIGAR over q and t (both fully consumed): the independent check of a path."""
    import re
    table = np.asarray(table).reshape(32, 32)
    i = j = 0; s = 0
    for n, op in re.findall(r"(\d+)([MID])", cigar):
        n = int(n)
        if op == "M":
            for k in range(n):
                s += int(table[q[i + k], t[j + k]])
            i += n; j += n
        elif op == "I":
            s += gap_open + n * gap_extend; i += n
        else:
            s += gap_open + n * gap_extend; j += n
    assert i == len(q) and j == len(t), (i, j, len(q), len(t))
    return s


def search32(q, codes, off, table, gap_open, gap_extend, threads=0):
    """int32 scalar oracle over a whole database (OpenMP). Returns int32[n]."""
    q, codes, off, table = _prep(q, codes, off, table)
    n = off.size - 1
    out = np.zeros(n, dtype=np.int32)
    oracle().swo_search32(_p(q), q.size, _p(codes), _p(off), n, _p(table), gap_open, gap_extend, threads, _p(out))
    return out


def ref_search16(q, codes, off, table, gap_open, gap_extend, threads=0):
    """The reference's own AVX2 kernel (libalign) over a length-sorted database. Returns (int16[n], seconds)."""
    q, codes, off, table = _prep(q, codes, off, table)
    n = off.size - 1
    out = np.zeros(n, dtype=np.int16)
    secs = ref().refh_search16(_p(q), q.size, _p(codes), _p(off), n, _p(table), gap_open, gap_extend, threads, _p(out))
    return out, float(secs)
