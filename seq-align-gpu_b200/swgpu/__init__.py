"""swgpu -- ctypes binding of libswgpu.so (the C-ABI in include/swgpu.h).

Python is only the test / bench harness language here; the product is the C-ABI
library and the C host layer.  The binding refuses to work without the compiled
CUDA library: there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# SWGPU_LIB: profiling knob (scripts/ab_lib.py compares two builds of the same library on one box)
LIB_PATH = os.environ.get("SWGPU_LIB") or os.path.join(os.path.dirname(_HERE), "libswgpu.so")


class SwgpuError(RuntimeError):
    pass


class Stats(C.Structure):
    _fields_ = [
        ("ms_total", C.c_double), ("ms_inter", C.c_double), ("ms_intra", C.c_double),
        ("ms_rescore", C.c_double), ("ms_h2d", C.c_double), ("ms_d2h", C.c_double),
        ("cells", C.c_int64), ("cells_inter", C.c_int64), ("cells_intra", C.c_int64),
        ("cells_rescore", C.c_int64), ("n_rescored", C.c_int64), ("n_long", C.c_int64),
        ("kernel_launches", C.c_int32), ("tile_rows", C.c_int32),
        ("ms_topk", C.c_double), ("topk_launches", C.c_int32), ("lanes_per_pair", C.c_int32),
        ("passes", C.c_int32), ("reserved", C.c_int32),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class Hit(C.Structure):
    _fields_ = [("score", C.c_int32), ("q_start", C.c_int32), ("q_end", C.c_int32),
                ("t_start", C.c_int32), ("t_end", C.c_int32), ("entry", C.c_int64)]


class BatchStats(C.Structure):
    _fields_ = [("ms_total", C.c_double), ("cells", C.c_int64), ("n_rescored", C.c_int64),
                ("kernel_launches", C.c_int32), ("slots", C.c_int32), ("reserved", C.c_int32 * 4)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ if k != "reserved"}


class SynthGen(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("n_letters", C.c_int32), ("reserved", C.c_int32),
                ("cdf", C.c_uint32 * 32), ("codes", C.c_int8 * 32)]


class PackOpts(C.Structure):
    _fields_ = [("long_threshold", C.c_int32), ("stream_cols", C.c_int32), ("no_taper", C.c_int32),
                ("reserved", C.c_int32 * 5)]


_lib = None


def lib():
    """Load libswgpu.so; raise loudly if the CUDA extension was not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise SwgpuError(f"{LIB_PATH} missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                         "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    P = C.POINTER
    L.swgpu_device_count.restype = C.c_int
    L.swgpu_last_error.restype = C.c_char_p
    L.swgpu_version.restype = C.c_char_p
    L.swgpu_db_create.restype = C.c_int
    L.swgpu_db_create.argtypes = [P(C.c_void_p), C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p]
    L.swgpu_db_create_synthetic.restype = C.c_int
    L.swgpu_db_create_synthetic.argtypes = [P(C.c_void_p), C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64,
                                            P(SynthGen), C.c_void_p]
    L.swgpu_db_fetch_target.restype = C.c_int
    L.swgpu_db_fetch_target.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, P(C.c_int64)]
    L.swgpu_db_destroy.argtypes = [C.c_void_p]
    L.swgpu_db_destroy.restype = None
    for f in ("swgpu_db_num_targets", "swgpu_db_num_residues", "swgpu_db_device_bytes"):
        getattr(L, f).restype = C.c_int64
        getattr(L, f).argtypes = [C.c_void_p]
    L.swgpu_search.restype = C.c_int
    L.swgpu_search.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.c_int, C.c_void_p, P(Stats)]
    L.swgpu_search_resident.restype = C.c_int
    L.swgpu_search_resident.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.c_int, P(Stats)]
    L.swgpu_db_scores_d2h.restype = C.c_int
    L.swgpu_db_scores_d2h.argtypes = [C.c_void_p, C.c_void_p]
    L.swgpu_topk.restype = C.c_int
    L.swgpu_topk.argtypes = [C.c_void_p, C.c_int, P(Hit), P(C.c_int)]
    L.swgpu_search_many.restype = C.c_int
    L.swgpu_search_many.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                    C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, P(BatchStats)]
    L.swgpu_traceback.restype = C.c_int
    L.swgpu_traceback.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.c_int, P(Hit), C.c_int32,
                                  C.c_void_p, C.c_int64, C.c_void_p]
    L.swgpu_db_last_stats.restype = C.c_int
    L.swgpu_db_last_stats.argtypes = [C.c_void_p, P(Stats)]
    L.swgpu_fill_batch16.restype = C.c_int
    L.swgpu_fill_batch16.argtypes = [C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int,
                                     C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    _lib = L
    return L


def _check(rc):
    if rc != 0:
        raise SwgpuError(f"swgpu error {rc}: {lib().swgpu_last_error().decode()}")


def _i8(a):
    a = np.ascontiguousarray(a, dtype=np.int8)
    return a, a.ctypes.data_as(C.c_void_p)


class Database:
    """Packed, HBM-resident database on one GPU (swgpu_db_t)."""

    def __init__(self, codes, offsets, device=0, entry_base=0, long_threshold=0, stream_cols=0, no_taper=0):
        L = lib()
        self._codes, pc = _i8(codes)
        self._off = np.ascontiguousarray(offsets, dtype=np.int64)
        self.n = len(self._off) - 1
        h = C.c_void_p()
        opts = PackOpts(long_threshold=long_threshold, stream_cols=stream_cols, no_taper=no_taper)
        _check(L.swgpu_db_create(C.byref(h), device, pc, self._off.ctypes.data_as(C.c_void_p), self.n, entry_base,
                                 C.byref(opts)))
        self._h = h

    @classmethod
    def synthetic(cls, lengths, seed, entry_ids=None, entry_base=0, spec=None, device=0, long_threshold=0, stream_cols=0):
        """swgpu_db_create_synthetic(): the residues are generated on the device from (seed, global entry id, position);
        swgpu.synth.synth_target() is the host mirror."""
        from . import synth
        cdf, cd, nl = spec or synth.synth_spec()
        self = cls.__new__(cls)
        self._lens = np.ascontiguousarray(lengths, dtype=np.int32)
        self._ids = None if entry_ids is None else np.ascontiguousarray(entry_ids, dtype=np.int64)
        self.n = len(self._lens)
        gen = SynthGen(seed=int(seed), n_letters=int(nl))
        for i in range(32):
            gen.cdf[i] = int(cdf[i]); gen.codes[i] = int(cd[i])
        opts = PackOpts(long_threshold=long_threshold, stream_cols=stream_cols)
        h = C.c_void_p()
        _check(lib().swgpu_db_create_synthetic(C.byref(h), device, self._lens.ctypes.data_as(C.c_void_p),
                                               None if self._ids is None else self._ids.ctypes.data_as(C.c_void_p),
                                               self.n, int(entry_base), C.byref(gen), C.byref(opts)))
        self._h = h
        return self

    def fetch_target(self, local_entry):
        """Residues of one target as they sit in the packed database on the device."""
        ln = C.c_int64(0)
        _check(lib().swgpu_db_fetch_target(self._h, int(local_entry), None, 0, C.byref(ln)))
        out = np.empty(ln.value, dtype=np.int8)
        _check(lib().swgpu_db_fetch_target(self._h, int(local_entry), out.ctypes.data_as(C.c_void_p), out.size, C.byref(ln)))
        return out

    @property
    def n_residues(self):
        return lib().swgpu_db_num_residues(self._h)

    def close(self):
        if getattr(self, "_h", None):
            lib().swgpu_db_destroy(self._h)
            self._h = None

    __del__ = close

    @property
    def device_bytes(self):
        return lib().swgpu_db_device_bytes(self._h)

    def search(self, query, table, gap_open, gap_extend, out=None):
        """Host API: H2D query/table, kernels, D2H scores (file order). Returns (scores int32[n], stats dict)."""
        q, pq = _i8(query)
        t, pt = _i8(np.asarray(table).reshape(-1))
        assert t.size == 1024
        scores = out if out is not None else np.empty(self.n, dtype=np.int32)
        st = Stats()
        _check(lib().swgpu_search(self._h, pq, q.size, pt, int(gap_open), int(gap_extend),
                                  scores.ctypes.data_as(C.c_void_p), C.byref(st)))
        return scores, st.as_dict()

    def search_resident(self, query, table, gap_open, gap_extend):
        q, pq = _i8(query)
        t, pt = _i8(np.asarray(table).reshape(-1))
        st = Stats()
        _check(lib().swgpu_search_resident(self._h, pq, q.size, pt, int(gap_open), int(gap_extend), C.byref(st)))
        return st.as_dict()

    def search_many(self, queries, table, gap_open, gap_extend, k=0, want_scores=True, out=None):
        """swgpu_search_many(): several queries in flight.  Returns (scores [nq][n] or None, hits per query,
        per-query stats, batch stats).  `out`: optional list of int32[n] host arrays (pinned ones are written
        directly by the copy engine)."""
        nq = len(queries)
        qs = [np.ascontiguousarray(q, dtype=np.int8) for q in queries]
        qptr = (C.c_void_p * max(nq, 1))(*[q.ctypes.data for q in qs])
        qlen = np.array([q.size for q in qs], dtype=np.int64)
        t, pt = _i8(np.asarray(table).reshape(-1))
        assert t.size == 1024
        scores, sptr = None, None
        if want_scores:
            scores = out if out is not None else [np.empty(self.n, dtype=np.int32) for _ in range(nq)]
            sptr = (C.c_void_p * max(nq, 1))(*[a.ctypes.data for a in scores])
        hits = (Hit * max(nq * k, 1))()
        nh = np.zeros(max(nq, 1), dtype=np.int32)
        per = (Stats * max(nq, 1))()
        bs = BatchStats()
        _check(lib().swgpu_search_many(self._h, nq, qptr, qlen.ctypes.data_as(C.c_void_p), pt, int(gap_open),
                                       int(gap_extend), sptr, int(k), hits if k else None,
                                       nh.ctypes.data_as(C.c_void_p) if k else None, per, C.byref(bs)))
        out_hits = [[dict(score=h.score, q_start=h.q_start, q_end=h.q_end, t_start=h.t_start, t_end=h.t_end,
                          entry=h.entry) for h in hits[q * k: q * k + int(nh[q])]] for q in range(nq)] if k else None
        return scores, out_hits, [per[q].as_dict() for q in range(nq)], bs.as_dict()

    def traceback(self, query, table, gap_open, gap_extend, hits):
        """swgpu_traceback(): CIGAR strings ('M' aligned pair, 'I' query residue vs gap, 'D' target residue vs gap)
        for a list of hit dicts as returned by topk() / search_many()."""
        q, pq = _i8(query)
        t, pt = _i8(np.asarray(table).reshape(-1))
        n = len(hits)
        arr = (Hit * max(n, 1))()
        cap = 0
        for i, h in enumerate(hits):
            arr[i] = Hit(h["score"], h["q_start"], h["q_end"], h["t_start"], h["t_end"], h["entry"])
            if h["q_start"] >= 0:
                cap += (h["q_end"] - h["q_start"] + 1) + (h["t_end"] - h["t_start"] + 1)
        ops = np.zeros(max(cap, 1), dtype=np.uint32)
        off = np.zeros(n + 1, dtype=np.int64)
        _check(lib().swgpu_traceback(self._h, pq, q.size, pt, int(gap_open), int(gap_extend), arr, n,
                                     ops.ctypes.data_as(C.c_void_p), ops.size, off.ctypes.data_as(C.c_void_p)))
        out = []
        for i in range(n):
            o = ops[off[i]:off[i + 1]]
            out.append("".join(f"{int(x) >> 2}{'MID'[int(x) & 3]}" for x in o))
        return out

    def scores(self):
        s = np.empty(self.n, dtype=np.int32)
        _check(lib().swgpu_db_scores_d2h(self._h, s.ctypes.data_as(C.c_void_p)))
        return s

    def last_stats(self):
        st = Stats()
        _check(lib().swgpu_db_last_stats(self._h, C.byref(st)))
        return st.as_dict()

    def topk(self, k):
        hits = (Hit * k)()
        n = C.c_int(0)
        _check(lib().swgpu_topk(self._h, k, hits, C.byref(n)))
        st = self.last_stats()
        self.last_topk_ms, self.last_topk_launches = st["ms_topk"], st["topk_launches"]
        return [dict(score=h.score, q_start=h.q_start, q_end=h.q_end, t_start=h.t_start, t_end=h.t_end, entry=h.entry)
                for h in hits[: n.value]]


def fill_batch16(query, batch_interleaved, padded_len, lanes, table, gap_open, gap_extend, device=0):
    q, pq = _i8(query)
    b, pb = _i8(batch_interleaved)
    t, pt = _i8(np.asarray(table).reshape(-1))
    out = np.zeros(16, dtype=np.int32)
    _check(lib().swgpu_fill_batch16(device, pq, q.size, pb, padded_len, lanes, pt, gap_open, gap_extend,
                                    out.ctypes.data_as(C.c_void_p)))
    return out[:lanes]
