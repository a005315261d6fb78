"""GPU suite (-m gpu), collected LAST on purpose: swgpu_score_targets_noalloc(), the entry
point that allocates nothing (include/swgpu.h; no counterpart in the reference).
Its scratch memory is a file-scope static __device__ array shared by every call, so the test calls it
twice in a row (different queries: one stripe, then several stripes through the carry rows) through the C-ABI and
compares each result with the existing entry point swgpu_search() on the same inputs -- integer work, bit-exact.
Passes alone in a fresh process and after the other tests (it keeps no state but the library's own)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_LEN = 4096            # SWGPU_NOALLOC_MAX_TARGET_LEN


def entry(swgpu):
    L = swgpu.lib()
    f = L.swgpu_score_targets_noalloc
    f.restype = C.c_int
    f.argtypes = [C.c_int, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int, C.c_int,
                  C.c_void_p, C.c_void_p]
    return L, f


def test_noalloc_entry_point_equals_swgpu_search_on_two_calls_in_a_row(matrices):
    import torch
    import swgpu
    from swgpu import synth
    assert torch.cuda.is_available() and swgpu.lib().swgpu_device_count() >= 1, "needs a CUDA device; there is no CPU fallback"
    L, f = entry(swgpu)
    tab = np.ascontiguousarray(matrices["BLOSUM62"], dtype=np.int8)
    # 700 targets: log-normal lengths, two empty ones, one at exactly the limit, some related to the queries
    g = synth.rng(2718)
    lens = synth.lognormal_lengths(g, 700, lo=1, hi=3000)
    lens[5] = 0; lens[333] = 0; lens[17] = MAX_LEN
    off = np.zeros(len(lens) + 1, dtype=np.int64); np.cumsum(lens, out=off[1:])
    codes = synth.residues(g, off[-1])
    q_long, q_short = synth.make_query(31, 700), synth.make_query(32, 100)      # 3 stripes of 256 rows; one stripe of 128
    codes[off[17] + 500: off[17] + 1100] = q_long[50:650]
    codes[off[40]: off[40] + min(int(lens[40]), 90)] = q_short[:min(int(lens[40]), 90)]
    db = swgpu.Database(codes, off)                                              # the existing entry point
    dev = torch.device("cuda", 0)
    d_codes = torch.from_numpy(codes).to(dev)
    d_off = torch.from_numpy(off).to(dev)
    d_tab = torch.from_numpy(tab.reshape(-1)).to(dev)
    d_scores = torch.full((len(lens),), -7, dtype=torch.int32, device=dev)
    for call, (q, go, ge) in enumerate([(q_long, -10, -1), (q_short, -10, -1), (q_long, -3, -2)]):
        want, _ = db.search(q, tab, go, ge)
        d_q = torch.from_numpy(np.ascontiguousarray(q)).to(dev)
        d_scores.fill_(-7)
        torch.cuda.synchronize()
        rc = f(0, d_q.data_ptr(), len(q), d_codes.data_ptr(), d_off.data_ptr(), len(lens), MAX_LEN, d_tab.data_ptr(), go, ge,
               d_scores.data_ptr(), None)
        assert rc == 0, L.swgpu_last_error().decode()
        torch.cuda.synchronize()
        got = d_scores.cpu().numpy()
        bad = np.nonzero(got != want)[0]
        assert bad.size == 0, (call, bad[:8], got[bad[:8]], want[bad[:8]])
    assert int(want[17]) > 500                                                   # the planted segment was found
    # on a stream of the caller's, twice back to back without a synchronisation in between (stream order serialises them)
    s = torch.cuda.Stream(device=dev)
    d_q = torch.from_numpy(np.ascontiguousarray(q_long)).to(dev)
    d_a = torch.empty(len(lens), dtype=torch.int32, device=dev); d_b = torch.empty_like(d_a)
    torch.cuda.synchronize()
    for out in (d_a, d_b):
        assert f(0, d_q.data_ptr(), len(q_long), d_codes.data_ptr(), d_off.data_ptr(), len(lens), MAX_LEN, d_tab.data_ptr(), -10, -1,
                 out.data_ptr(), C.c_void_p(s.cuda_stream)) == 0, L.swgpu_last_error().decode()
    s.synchronize()
    want, _ = db.search(q_long, tab, -10, -1)
    assert np.array_equal(d_a.cpu().numpy(), want) and np.array_equal(d_b.cpu().numpy(), want)
    db.close()


def test_noalloc_entry_point_rejects_sizes_beyond_the_static_scratch():
    import swgpu
    L, f = entry(swgpu)
    rc = f(0, None, 10, None, None, 4, MAX_LEN + 1, None, -10, -1, None, None)
    assert rc == -3 and "SWGPU_NOALLOC_MAX_TARGET_LEN" in L.swgpu_last_error().decode()
    assert f(0, None, 10, None, None, -1, 100, None, -10, -1, None, None) == -3
    assert f(0, None, 10, None, None, 4, 100, None, 5, -1, None, None) == -3         # a positive gap score: this entry point has the two-gap-state kernel only
