"""CPU suite: host logic of the device-generated databases (swgpu_db_create_synthetic).

  * db_pack.c::swpack_layout (layout from lengths alone) produces exactly the tables swpack_build produces from real
    residues, and its pair_col[] points at the flagged first column of every pair in swpack_build's streams;
  * the host mirror of the generator (swgpu.synth.synth_target) is deterministic, depends on (seed, entry, position)
    only, and draws the Swiss-Prot background composition;
  * the C5 workload definition (lengths, planted queries) has the shape SURVEY.md 8(d) names;
  * swgpu_db_create_synthetic fails loudly without a CUDA device (no CPU fallback).
No GPU work is launched."""
import ctypes as C

import numpy as np
import pytest


class SwPack(C.Structure):
    _fields_ = [("n_targets", C.c_int64), ("n_residues", C.c_int64), ("max_len", C.c_int32),
                ("order", C.POINTER(C.c_int64)), ("plain_off", C.POINTER(C.c_int64)), ("plain", C.POINTER(C.c_int8)),
                ("n_pairs", C.c_int64), ("n_streams", C.c_int64), ("max_stream_cols", C.c_int32),
                ("n_long_streams", C.c_int64), ("n_long_pairs", C.c_int64),
                ("scols", C.POINTER(C.c_uint16)), ("scols_len", C.c_int64),
                ("stream_off", C.POINTER(C.c_int64)), ("stream_ncols", C.POINTER(C.c_int32)),
                ("stream_pair0", C.POINTER(C.c_int64))]


def arr(ptr, n, dtype):
    return np.ctypeslib.as_array(ptr, shape=(max(int(n), 1),))[: int(n)].astype(dtype).copy()


@pytest.mark.parametrize("taper", [0, 3])
@pytest.mark.parametrize("seed,n,stream_cols", [(1, 1, 1536), (2, 2, 1536), (3, 257, 64), (4, 5000, 1536), (5, 4001, 300)])
def test_layout_from_lengths_matches_the_packer(seed, n, stream_cols, taper):
    import swgpu
    from swgpu import synth
    L = swgpu.lib()
    g = synth.rng(seed)
    lens = synth.lognormal_lengths(g, n, median=230.0, sigma=0.75, lo=0 if seed == 3 else 11, hi=35213).astype(np.int32)
    if n > 100:
        lens[7] = 9000; lens[8] = 0                           # a long stream of its own, an empty target
    off = np.zeros(n + 1, dtype=np.int64); np.cumsum(lens, out=off[1:])
    codes = synth.residues(g, int(off[-1]))
    a, b = SwPack(), SwPack()
    assert L.swpack_build(C.byref(a), codes.ctypes.data_as(C.c_void_p), off.ctypes.data_as(C.c_void_p), C.c_int64(n),
                          C.c_int32(stream_cols), C.c_int32(taper)) == 0
    pair_col = np.zeros((n + 1) // 2 + 1, dtype=np.int64)
    assert L.swpack_layout(C.byref(b), lens.ctypes.data_as(C.c_void_p), C.c_int64(n), C.c_int32(stream_cols),
                           C.c_int32(taper), pair_col.ctypes.data_as(C.c_void_p)) == 0
    try:
        for f in ("n_targets", "n_residues", "max_len", "n_pairs", "n_streams", "max_stream_cols", "n_long_streams",
                  "n_long_pairs", "scols_len"):
            assert getattr(a, f) == getattr(b, f), f
        assert not b.plain and not b.scols                     # layout only
        assert np.array_equal(arr(a.order, n, np.int64), arr(b.order, n, np.int64))
        assert np.array_equal(arr(a.plain_off, n + 1, np.int64), arr(b.plain_off, n + 1, np.int64))
        ns = a.n_streams
        assert np.array_equal(arr(a.stream_off, ns + 1, np.int64), arr(b.stream_off, ns + 1, np.int64))
        assert np.array_equal(arr(a.stream_ncols, ns, np.int32), arr(b.stream_ncols, ns, np.int32))
        assert np.array_equal(arr(a.stream_pair0, ns, np.int64), arr(b.stream_pair0, ns, np.int64))
        # pair_col: the flagged columns of the packer's streams, in pair order
        scols = arr(a.scols, a.scols_len, np.uint16)
        flagged = np.flatnonzero(scols & 0x80)
        assert np.array_equal(flagged, pair_col[: a.n_pairs])
        # and the pair's columns there are its two members, as the device kernel will write them
        order, poff, plain = arr(a.order, n, np.int64), arr(a.plain_off, n + 1, np.int64), arr(a.plain, a.n_residues, np.int8)
        for q in range(0, int(a.n_pairs), max(1, int(a.n_pairs) // 50)):
            la = int(poff[2 * q + 1] - poff[2 * q]); lb = int(poff[2 * q + 2] - poff[2 * q + 1]) if 2 * q + 1 < n else 0
            col = scols[pair_col[q]: pair_col[q] + max(la, 1)].astype(np.int64)
            lo = np.zeros(max(la, 1), dtype=np.int64); lo[:la] = plain[poff[2 * q]: poff[2 * q] + la]
            hi = np.zeros(max(la, 1), dtype=np.int64); hi[:lb] = plain[poff[2 * q + 1]: poff[2 * q + 1] + lb]
            want = lo | (hi << 8); want[0] |= 0x80
            assert np.array_equal(col, want), q
        # taper (number of halvings): full-length streams first, half-length ones in the last 1/16 of the columns,
        # quarter-length in the last 1/64, an eighth in the last 1/256
        ncols, soff = arr(a.stream_ncols, ns, np.int32), arr(a.stream_pair0, ns, np.int64)
        if n >= 4000 and stream_cols == 1536:
            body = ncols[1: ns // 2]                               # (stream 0 may be the 9 000-column pair)
            assert body.min() >= stream_cols
            if taper:
                assert ncols[-2] < stream_cols // 8 + 400
                assert np.any((ncols >= stream_cols // 2) & (ncols < stream_cols)) and np.any((ncols >= stream_cols // 4) & (ncols < stream_cols // 2))
            else:
                assert ncols[:-1].min() >= stream_cols
    finally:
        L.swpack_free(C.byref(a)); L.swpack_free(C.byref(b))


def test_generator_mirror_is_a_pure_function_of_seed_entry_position():
    from swgpu import synth
    spec = synth.synth_spec()
    cdf, cd, nl = spec
    assert nl == 20 and cdf[nl - 1] == 65536 and np.all(np.diff(cdf[:nl].astype(np.int64)) > 0)
    a = synth.synth_target(46, 12345, 500, spec)
    assert np.array_equal(a, synth.synth_target(46, 12345, 500, spec))
    assert np.array_equal(a[:200], synth.synth_target(46, 12345, 200, spec))          # prefix property
    assert not np.array_equal(a, synth.synth_target(46, 12346, 500, spec))
    assert not np.array_equal(a, synth.synth_target(47, 12345, 500, spec))
    assert set(np.unique(a)) <= set(int(c) for c in synth.CODES)
    # known answer, frozen, and the formula of include/swgpu.h restated with Python integers
    assert synth.synth_target(46, 0, 12, spec).tolist() == [1, 4, 12, 1, 16, 22, 20, 19, 19, 5, 17, 19]
    for seed, e, ln in ((46, 0, 40), (46, 32_797_638, 64), (2**63 + 11, 2**40 + 7, 33)):
        assert synth.synth_target(seed, e, ln, spec).tolist() == plain_python_target(seed, e, ln, spec)
    big = synth.synth_target(1, 2**40 + 7, 400_000, spec)
    freq = np.array([(big == c).mean() for c in synth.CODES])
    assert np.abs(freq - synth.FREQ).max() < 0.003


def plain_python_target(seed, entry, length, spec):
    cdf, cd, nl = spec
    M = (1 << 64) - 1
    out = []
    for j in range(length):
        x = ((seed ^ (entry * 0x9E3779B97F4A7C15 & M)) + (j + 1) * 0xBF58476D1CE4E5B9) & M
        x ^= x >> 30; x = x * 0xBF58476D1CE4E5B9 & M
        x ^= x >> 27; x = x * 0x94D049BB133111EB & M
        x ^= x >> 31
        u = x >> 48
        i = 0
        while i < nl - 1 and u >= int(cdf[i]):
            i += 1
        out.append(int(cd[i]))
    return out


def test_c5_workload_shape():
    from swgpu import synth
    lens = synth.c5_lengths(0.001)
    assert lens.dtype == np.int32 and lens.min() >= 11 and lens.max() == 35213 and lens[0] == 35213
    assert abs(lens.astype(np.int64).sum() / (synth.C5_TOTAL_RESIDUES * 0.001) - 1) < 0.02
    assert abs(lens.mean() - 305) < 8
    qs, planted = synth.c5_queries(lens, 64)
    assert [len(q) for q in qs] == [synth.C2_QUERY_LENGTHS[i % 20] for i in range(64)]
    assert len(set(planted)) == 64 and all(lens[e] >= 200 for e in planted)
    # the planted segment is really there: >= 80 % identity with the target's head somewhere in the query
    for i in (0, 9, 37):
        t = synth.synth_target(synth.C5_SEED, planted[i], int(lens[planted[i]]))
        m = min(len(qs[i]), len(t))
        best = max(float((qs[i][a:a + m] == t[:m]).mean()) for a in range(len(qs[i]) - m + 1))
        assert best > 0.8, (i, best)


def test_synthetic_create_fails_loudly_without_a_device():
    import swgpu
    if swgpu.lib().swgpu_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(swgpu.SwgpuError) as e:
        swgpu.Database.synthetic(np.array([10, 20], dtype=np.int32), 46)
    assert "swgpu error -1" in str(e.value) or "swgpu error -2" in str(e.value)


def test_c5_bench_shards_partition_the_database():
    """bench.py --workload c5 (and c2 --sharded) under N ranks: the shards are a partition of the global entry ids, balanced in residues
    (within one target) and in length mix; the queries do not depend on the rank."""
    import importlib.util, os
    from conftest import ROOT
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec); spec.loader.exec_module(bench)
    import argparse
    world = 4

    def load(name, rank, w, **kw):
        a = argparse.Namespace(workload=name, scale=0.0005, sharded=False, shards_total=8)
        for k, v in kw.items():
            setattr(a, k, v)
        d = bench.load_workload(a, rank, w)
        return d["lens"], d["entries"], d["queries"]

    for name, kw in (("c5", {}), ("c2", {"sharded": True, "shards_total": 3, "scale": 0.002})):
        parts = [load(name, r, world, **kw) for r in range(world)]
        if name == "c5":
            full_lens, none_entries, q1 = load(name, 0, 1, **kw)
            assert none_entries is None
        else:                                                    # --sharded deals even a single rank by global entry id
            full_lens = bench.c2_sharded_lengths(3, 0.002)
            l1, e1, q1 = load(name, 0, 1, **kw)
            assert np.array_equal(e1, np.arange(len(full_lens))) and np.array_equal(l1, full_lens)
        ids = np.concatenate([p[1] for p in parts])
        assert np.array_equal(np.sort(ids), np.arange(len(full_lens)))
        for lens, entries, qs in parts:
            assert np.array_equal(lens, full_lens[entries]) and np.all(np.diff(entries) > 0)
            assert all(np.array_equal(a, b) for a, b in zip(qs, q1))
        res = [int(p[0].astype(np.int64).sum()) for p in parts]
        assert max(res) - min(res) <= 35213
        assert max(int(p[0].max()) for p in parts[1:]) >= np.sort(full_lens)[-world]


@pytest.mark.parametrize("seed", range(12))
def test_packer_invariants_on_random_databases(seed):
    """db_pack.c on random databases (empty targets, one giant, odd counts, any stream length and taper depth):
    every residue of every target sits exactly once in `plain` (sorted by length, stable) and once in the streams, in
    the half of its pair; pairs never straddle streams; streams start 64-column aligned; the long streams are a prefix."""
    import swgpu
    from swgpu import synth
    L = swgpu.lib()
    g = synth.rng(1000 + seed)
    n = int(g.integers(1, 700))
    lens = np.where(g.random(n) < 0.1, 0, g.integers(1, 400, size=n)).astype(np.int64)
    if seed % 3 == 0:
        lens[int(g.integers(0, n))] = int(g.integers(4097, 9000))
    off = np.zeros(n + 1, dtype=np.int64); np.cumsum(lens, out=off[1:])
    codes = g.integers(1, 32, size=int(off[-1])).astype(np.int8)
    stream_cols, taper = int(g.choice([32, 64, 100, 777, 1536, 2304])), int(g.integers(0, 5))
    p = SwPack()
    assert L.swpack_build(C.byref(p), codes.ctypes.data_as(C.c_void_p), off.ctypes.data_as(C.c_void_p), C.c_int64(n),
                          C.c_int32(stream_cols), C.c_int32(taper)) == 0
    try:
        order, poff = arr(p.order, n, np.int64), arr(p.plain_off, n + 1, np.int64)
        plain = arr(p.plain, p.n_residues, np.int8)
        assert sorted(order.tolist()) == list(range(n))
        sl = lens[order]
        assert np.all(np.diff(sl) <= 0)                                            # longest first ...
        assert all(order[i] < order[i + 1] for i in range(n - 1) if sl[i] == sl[i + 1])   # ... stable
        assert np.array_equal(np.diff(poff), sl) and p.n_residues == off[-1] and p.max_len == lens.max()
        for s in range(n):
            assert np.array_equal(plain[poff[s]:poff[s + 1]], codes[off[order[s]]:off[order[s] + 1]])
        ns, npairs = int(p.n_streams), int(p.n_pairs)
        assert npairs == (n + 1) // 2
        soff, ncols, pair0 = arr(p.stream_off, ns + 1, np.int64), arr(p.stream_ncols, ns, np.int32), arr(p.stream_pair0, ns, np.int64)
        scols = arr(p.scols, p.scols_len, np.uint16)
        assert np.all(soff % 64 == 0) and pair0[0] == 0 and np.all(np.diff(pair0) > 0) and p.max_stream_cols == ncols.max()
        nl = int(p.n_long_streams)
        first_pair = np.array([int(sl[2 * q0]) for q0 in pair0])              # a long stream = its FIRST pair alone exceeds 4096 columns
        assert np.all(first_pair[:nl] > 4096) and (nl == ns or first_pair[nl] <= 4096), (nl, first_pair[:nl + 1])
        q = 0
        for st in range(ns):
            col = soff[st]
            q_end = pair0[st + 1] if st + 1 < ns else npairs
            assert q == pair0[st]
            while q < q_end:
                la = int(sl[2 * q]); lb = int(sl[2 * q + 1]) if 2 * q + 1 < n else 0
                w = max(la, 1)
                seg = scols[col: col + w].astype(np.int64)
                assert seg[0] & 0x80 and not np.any(seg[1:] & 0x80)                # one flag, on the pair's first column
                lo, hi = seg & 0x1f, (seg >> 8) & 0x1f
                assert np.array_equal(lo[:la], plain[poff[2 * q]: poff[2 * q] + la]) and not np.any(lo[la:])
                assert np.array_equal(hi[:lb], plain[poff[2 * q + 1]: poff[2 * q + 1] + lb] if lb else hi[:0]) and not np.any(hi[lb:])
                col += w; q += 1
            assert col - soff[st] == ncols[st] and col <= soff[st + 1]
            assert not np.any(scols[col: soff[st + 1]])                            # alignment padding is code 0
        assert q == npairs
    finally:
        L.swpack_free(C.byref(p))
