"""GPU suite: the K1S planner on databases that are NOT the benchmark's.  Its constants (rows-per-lane penalties, the
multi-stripe cost table, swgpu.cu plan_k1s) were measured on the C2 database; this guards the choice elsewhere: on two
differently shaped databases -- few, long targets (median 1 500 aa, 12 k of them) and many short ones (median 120 aa,
200 k) -- the plan the planner picks must be within 12 % of the fastest admissible plan (every admissible (lanes, rows)
plan is forced through SWGPU_K1S_PLAN and timed, min of 3), for short, single-stripe and multi-stripe query lengths.
Scores of every forced plan are compared with the default plan's (bit-exact), so this is a parity test as well."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MULTI_R = (16, 20, 24, 25, 26, 27, 28, 29, 30, 31, 32)


def admissible(lq):
    plans = set()
    for g in (8, 16, 32):
        R = max(-(-lq // g), 8)
        for d in (0, 1, 2):
            if R + d <= 32:
                plans.add((g, R + d))
    if lq > 1024:
        plans = {(32, R) for R in MULTI_R}
    return sorted(plans)


def timed(db, q, tab, reps=3):
    best, st = 1e9, None
    for _ in range(reps):
        st = db.search_resident(q, tab, -10, -1)
        best = min(best, st["ms_total"])
    return best, st


@pytest.mark.parametrize("shape", ["few_long", "many_short"])
def test_planner_choice_is_near_the_best_plan_on_other_databases(matrices, shape):
    import swgpu
    from swgpu import synth
    tab = matrices["BLOSUM62"]
    if shape == "few_long":
        codes, off = synth.make_db(901, 12000, median=1500.0, sigma=0.5, lo=200, hi=20000)
        lqs = (200, 700, 2600)
    else:
        codes, off = synth.make_db(902, 200000, median=120.0, sigma=0.6, lo=10, hi=3000)
        lqs = (96, 450, 1500)
    db = swgpu.Database(codes, off)
    report = []
    for lq in lqs:
        q = synth.make_query(903 + lq, lq)
        os.environ.pop("SWGPU_K1S_PLAN", None)
        db.search_resident(q, tab, -10, -1)                                   # warm-up (buffers grow on first use)
        t_def, st = timed(db, q, tab)
        want = db.scores().copy()
        default = (st["lanes_per_pair"], st["tile_rows"], st["passes"])
        times = {}
        try:
            for g, R in admissible(lq):
                os.environ["SWGPU_K1S_PLAN"] = f"{g},{R}"
                t, st = timed(db, q, tab)
                if (st["lanes_per_pair"], st["tile_rows"]) != (g, R):
                    continue                                                   # no such instantiation for this length
                assert np.array_equal(db.scores(), want), (shape, lq, g, R)
                times[(g, R, st["passes"])] = t
        finally:
            os.environ.pop("SWGPU_K1S_PLAN", None)
        best = min(times, key=times.get)
        report.append((lq, default, round(t_def, 3), best, round(times[best], 3)))
        assert t_def <= 1.12 * times[best] + 0.05, (shape, lq, default, t_def, best, times[best], times)
    print(f"planner on {shape}: " + "; ".join(f"Lq {lq}: {d} {td} ms (best {b} {tb} ms)" for lq, d, td, b, tb in report))
    db.close()
