#!/usr/bin/env python
"""bench.py -- GCUPS of the one-to-many Smith-Waterman scoring path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c1|c2|c3|c4|c5]
                  [--scale F] [--sharded [--shards-total M]] [--cpu-seconds S] [--no-cpu] [--no-parity]
  torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...        (N > 1: one rank per GPU, NCCL)

Workloads (BASELINE.json `configs`); the default, and the line the driver reads, is c2:
  c1  360-aa query x 10 000 synthetic proteins, BLOSUM62 (the reference's own CPU-runnable case); step = the one query
  c2  Swiss-Prot-shaped synthetic database (570 000 sequences, ~205 M residues, log-normal lengths, one 35 213-aa
      target; seed 43 + rank), BLOSUM62, the classic 20-query length sweep 144..5478 aa.  One step = the whole sweep
      over the HBM-resident database = 41 752 x n_residues cell updates per GPU.  N > 1 is weak scaling: every rank
      holds its own 205 M-residue shard (the database shards by residue count, no data-path collective) and the
      per-rank top-100 lists of the 20 queries are merged with ONE NCCL all_gather per step -- inside `value`.
      --sharded: ONE database of M x 205 M residues (default M = 8), generated on the device from (seed, global
      entry id), dealt longest-first round-robin over the ranks with global entry ids (strong scaling); the merged
      hit lists do not depend on N (`topk_crc`).
  c3  16-bit saturation stress: 4 096 targets related to a 6 000-aa low-complexity query, PAM250; step = the query
  c4  long sequences: 35 213-aa query x 1 024 targets up to 35 213 aa, BLOSUM62; step = the query
  c5  UniRef50-shaped database of 10 G residues (32.8 M targets) GENERATED ON THE DEVICE, dealt over the ranks
      (strong scaling), 64 queries cycling through the 20 lengths, BLOSUM45, top-100 per query.  --scale shrinks it.
All with gap open 10 / extend 1 (--gapopen -10 --gapextend -1 in the reference's convention).

value  = GCUPS, device-timed (CUDA events on the engine's streams: first kernel launch of the step -> last top-k
         result of the step in host staging; for N > 1 plus the CUDA-event time of the exchange H2D + all_gather +
         D2H and the host time of the merge), real unpadded cells of all ranks / max-over-ranks time.  The queries
         of a step go through swgpu_search_many(): up to three in flight, one query's tail overlaps the next.
e2e    = same metric through the host-facing C-ABI call swgpu_search_many() with HOST buffers: queries + table in
         host memory, H2D inside, kernels, D2H of all per-entry scores of every query into pinned host memory, host
         wall clock (no top-k: the caller gets every score, as the reference's caller does).
roofline = dominant kernel (K1S, packed 16-bit inter-sequence) against the measured integer-ALU-pipe cell-update
         roofline (DESIGN.md "Roofline"); the HBM view is reported beside it.
cpu_baseline = the reference's own AVX2/OpenMP kernel (oracle/_ref/libalign_ref.so, the unmodified reference
         sources) on the box's host cores, on a bounded sample of the same workload, plus a one-thread row.
parity = the scores of THIS run checked against the checker: the per-entry scores of the sampled batches (the
         same batches and queries the cpu_baseline leg computes with the reference's kernel) and every reported
         top hit (int32 scalar oracle).  A mismatch makes the run exit non-zero.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "seq-align-gpu_b200"))

GAP_OPEN, GAP_EXTEND = -10, -1
N_TARGETS, N_RESIDUES = 570_000, 205_000_000
TOPK = 100
C2_CPU_QUERIES = (0, 3, 9, 19)            # 144, 375, 1000, 5478 aa


def host_threads():
    """Host cores this process may use.  NOT omp_get_max_threads(): torchrun exports OMP_NUM_THREADS=1."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except (AttributeError, OSError):
        return max(1, os.cpu_count() or 1)


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d.get("hbm_gbs", 6650.0)), float(d.get("sm_max_mhz", 1965.0)), "measured"
    return 6650.0, 1965.0, "fallback"


def int_pipe_peak():
    """Measured ALU-pipe rate (warp-instructions / clk / SM for VIMNMX/VIADDMNMX.16x2), written by
    bench_micro/ubench_int on a B200 of this pool and committed as profiles/int_pipe_peak.json."""
    p = os.path.join(ROOT, "profiles", "int_pipe_peak.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["alu_winst_per_clk_sm"]), "measured (profiles/int_pipe_peak.json)"
    return 2.0, "nominal 64 lanes/clk/SM (no microbenchmark file)"


class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, pw, reasons = [], 0.0, [], set()
        for ts, line in self.rows:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 8 or not (t0 - 0.05 <= ts <= t1 + 0.15):
                continue
            try:
                sm.append(float(f[1])); mx = max(mx, float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


# ----------------------------------------------------------------------------------------------------------------
# workloads
# ----------------------------------------------------------------------------------------------------------------
def c2_sharded_lengths(shards_total, scale):
    """Lengths of ONE database made of `shards_total` C2-shaped parts (seeds 43, 44, ...), file order = part by part."""
    from swgpu import synth
    n = max(64, int(N_TARGETS * scale)); total = int(N_RESIDUES * scale)
    out = []
    for s in range(shards_total):
        g = synth.rng(43 + s)
        lens = synth.lognormal_lengths(g, n, lo=2, hi=35213)
        lens = np.clip(np.floor(lens * (total / lens.sum())), 2, 35213).astype(np.int64)
        if scale >= 0.01:
            lens[0] = 35213
        out.append(lens)
    return np.concatenate(out).astype(np.int32)


def load_workload(a, rank, world):
    """Returns a dict: name, config, matrix, tab, queries, topk, scaling, and either (codes, off) -- a host database --
    or (lens, entries, seed) -- a database generated on the device; cpu_queries = indices the CPU legs time."""
    from swgpu import synth
    name = a.workload
    w = {"name": name, "topk": TOPK, "scaling": "weak", "codes": None, "off": None, "lens": None, "entries": None}
    mat = {"c1": "BLOSUM62", "c2": "BLOSUM62", "c3": "PAM250", "c4": "BLOSUM62", "c5": "BLOSUM45"}[name]
    w["matrix"] = mat
    w["tab"], _ = synth.load_matrix(os.path.join(ROOT, "scoring", mat + ".txt"))
    cache = "database + scratch per GPU exceed L2 (126 MB); every query pass streams the whole DB"
    if name == "c2" and not a.sharded:
        n = max(64, int(N_TARGETS * a.scale))
        w["codes"], w["off"] = synth.make_db(43 + rank, n, lo=2, hi=35213, total=int(N_RESIDUES * a.scale),
                                             force_max=35213 if a.scale >= 0.01 else None)
        w["queries"] = [synth.make_query(100 + i, L) for i, L in enumerate(synth.C2_QUERY_LENGTHS)]
        w["cpu_queries"] = C2_CPU_QUERIES
        w["entry_base"] = rank * n
        w["config"] = {"workload": "C2: Swiss-Prot-shaped synthetic DB (570k seqs, ~205M residues per GPU), BLOSUM62, gap open 10 / "
                                   "extend 1, 20-query sweep 144..5478 aa; one step = the whole sweep",
                       "targets_per_gpu": n, "queries": 20, "topk": TOPK, "cache": cache,
                       "sharding": "one 205M-residue shard per rank, no data-path collective; ONE NCCL all_gather of the per-rank "
                                   "top-k lists per step, inside the timed window"}
    elif name == "c2":
        lens = c2_sharded_lengths(a.shards_total, a.scale)
        order = np.argsort(-lens.astype(np.int64), kind="stable")
        w["entries"] = np.sort(order[rank::world])
        w["lens"] = lens[w["entries"]]
        w["seed"] = 43
        w["queries"] = [synth.make_query(100 + i, L) for i, L in enumerate(synth.C2_QUERY_LENGTHS)]
        w["cpu_queries"] = ()
        w["scaling"] = "strong"
        w["config"] = {"workload": f"C2 --sharded: ONE Swiss-Prot-shaped synthetic DB of {a.shards_total} x 570k seqs / "
                                   f"{a.shards_total} x ~205M residues generated on the device, dealt over the ranks with global "
                                   "entry ids, BLOSUM62, gap open 10 / extend 1, 20-query sweep 144..5478 aa; one step = the sweep",
                       "targets_total": int(len(lens)), "queries": 20, "topk": TOPK, "cache": cache,
                       "sharding": "targets dealt longest-first round-robin over the ranks (total fixed: strong scaling), no "
                                   "data-path collective; ONE NCCL all_gather of the per-rank top-k lists per step"}
    elif name == "c5":
        lens = synth.c5_lengths(a.scale)
        w["queries"], _ = synth.c5_queries(lens, 64)
        if world > 1:
            order = np.argsort(-lens.astype(np.int64), kind="stable")
            w["entries"] = np.sort(order[rank::world])
            lens = lens[w["entries"]]
        w["lens"] = lens
        w["seed"] = synth.C5_SEED
        w["cpu_queries"] = ()
        w["scaling"] = "strong"
        w["config"] = {"workload": f"C5: UniRef50-shaped synthetic DB generated on the device ({a.scale:g} x 10 G residues in total, "
                                   "LogNormal(ln 230, 0.75) lengths), BLOSUM45, gap open 10 / extend 1, 64 queries cycling through "
                                   "144..5478 aa, top-100 per query; one step = the 64 queries",
                       "queries": 64, "topk": TOPK,
                       "cache": "the database per GPU exceeds L2 (126 MB) by orders of magnitude; every query pass streams it",
                       "sharding": "targets dealt longest-first round-robin over the ranks (total fixed: strong scaling), no data-path "
                                   "collective; ONE NCCL all_gather of the per-rank top-k lists per step"}
    else:
        if name == "c1":
            q, codes, off = synth.config1()
            text = "C1: one 360-aa query x 10 000 synthetic proteins (3.67 M residues), BLOSUM62, gap open 10 / extend 1; one step = the query"
        elif name == "c3":
            q, codes, off = synth.config3()
            text = ("C3: 16-bit saturation stress -- 4 096 targets (exact / mutated substrings of the query and background), "
                    "6 000-aa low-complexity query, PAM250, gap open 10 / extend 1; one step = the query")
        else:
            q, codes, off = synth.config4()
            text = ("C4: long sequences -- 35 213-aa query x 1 024 targets up to 35 213 aa (16 of them titin-length, one a "
                    "90 %-identity copy), BLOSUM62, gap open 10 / extend 1; one step = the query")
        w["codes"], w["off"], w["queries"], w["cpu_queries"] = codes, off, [q], (0,)
        w["entry_base"] = rank * (len(off) - 1)
        w["config"] = {"workload": text, "targets_per_gpu": int(len(off) - 1), "queries": 1, "topk": TOPK,
                       "cache": "L2 is flushed between steps by the engine's own scratch traffic only for the larger configs; "
                                "C1 (3.7 MB) is L2-resident by nature of the configuration",
                       "sharding": "every rank runs the same small database (replicas); the per-rank top-k lists are still merged"}
    return w


# ----------------------------------------------------------------------------------------------------------------
# CPU legs: the reference's own kernel on a bounded sample, and the parity check
# ----------------------------------------------------------------------------------------------------------------
def cpu_sample(codes, off, stride):
    """Every stride-th GROUP of 16 consecutive targets of the length-sorted database: the reference batches 16
    consecutive records and pads them to the first one (src/alignment_cmdline.c:429-459), so sampling whole
    groups keeps its batch composition (and its padding cost) exactly what it is on the full database.
    Returns (codes, offsets, entry indices)."""
    n = len(off) - 1
    groups = np.arange(stride // 2, n // 16, stride)        # mid-stride start: the one extreme batch (the 35 213-aa target,
                                                            # 16 lanes padded to it) is not over-weighted by a sparse sample
    idx = (groups[:, None] * 16 + np.arange(16)[None, :]).reshape(-1)
    lens = (off[1:] - off[:-1])[idx]
    o = np.zeros(len(idx) + 1, dtype=np.int64); np.cumsum(lens, out=o[1:])
    c = np.concatenate([codes[off[i]:off[i + 1]] for i in idx]) if len(idx) else np.zeros(0, np.int8)
    return c, o, idx


class CpuReference:
    """Times the reference's own kernel (unmodified sources, oracle/_ref) -- or, if that library did not travel, the C
    restatement -- on every stride-th batch of the workload's database, for the workload's CPU queries."""

    def __init__(self, w, threads=None):
        from oracle import oracle as O
        self.O, self.w = O, w
        self.use_ref = O.have_ref()
        self.threads = int(threads or host_threads())
        self.qs = [w["queries"][i] for i in w["cpu_queries"]]
        self.nb = (len(w["off"]) - 1) // 16
        self.stride = None

    def once(self, sample, threads):
        """One pass of the CPU queries over a sample.  Returns (seconds in the parallel region, cells, scores per query)."""
        c, o, _ = sample
        secs, cells, scores = 0.0, 0, []
        for q in self.qs:
            if self.use_ref:
                sc, s = self.O.ref_search16(q, c, o, self.w["tab"], GAP_OPEN, GAP_EXTEND, threads=threads)
            else:
                t0 = time.perf_counter()
                sc = self.O.search32(q, c, o, self.w["tab"], GAP_OPEN, GAP_EXTEND, threads=threads)
                s = time.perf_counter() - t0
            secs += s; cells += len(q) * int(o[-1]); scores.append(sc)
        return secs, cells, scores

    def calibrate(self, budget_s):
        """Pick the stride so that one pass takes about budget_s (stride 1 = the whole database)."""
        s0 = max(1, self.nb // 64)
        smp = cpu_sample(self.w["codes"], self.w["off"], s0)
        secs, cells, _ = self.once(smp, self.threads)           # also warms the thread pool
        full_cells = sum(len(q) for q in self.qs) * int(self.w["off"][-1])
        want = cells / max(secs, 1e-6) * budget_s
        self.stride = int(max(1, min(s0, round(full_cells / max(want, 1)))))
        self.sample = cpu_sample(self.w["codes"], self.w["off"], self.stride)
        return self.stride

    def run(self):
        secs, cells, scores = self.once(self.sample, self.threads)
        self.last_scores = scores
        return secs, cells

    def one_thread(self, budget_s=3.0):
        """The same kernel on ONE host thread, on a sparser sample (about budget_s)."""
        stride1 = int(min(max(self.nb, 1), max(self.stride, 1) * max(1, round(self.threads * 4.0 / budget_s))))
        smp = cpu_sample(self.w["codes"], self.w["off"], stride1)
        if len(smp[2]) == 0:
            smp = self.sample
            stride1 = self.stride
        secs, cells, _ = self.once(smp, 1)
        return {"value": cells / max(secs, 1e-9) / 1e9, "unit": "GCUPS", "cores": 1, "seconds": secs,
                "sample": f"every {stride1}-th batch ({len(smp[2])} targets, {int(smp[1][-1])} residues)"}

    def describe(self):
        c, o, idx = self.sample
        return (f"every {self.stride}-th batch of 16 consecutive targets (starting at batch {self.stride // 2}) of the length-sorted "
                f"{self.w['name'].upper()} database ({len(idx)} targets, {int(o[-1])} residues) x queries of "
                f"{', '.join(str(len(q)) for q in self.qs)} aa; timer = the reference's own region (omp parallel for over "
                f"alignment_fill_matrices, schedule(dynamic,1)); threads set with omp_set_num_threads({self.threads}) = the cores "
                f"this process may run on (OMP_NUM_THREADS is ignored: torchrun sets it to 1)")


def check_parity(w, cpu, gpu_scores, hit_lists, threads):
    """gpu_scores: {query index: int32[n] in file order}; hit_lists: per query [(score, local entry)].
    (1) the batches the CPU leg scored with the reference's kernel: equal wherever the reference's int16 result is
    exact (<= 32767 - max(matrix)), the others against the int32 scalar oracle; (2) every reported top
    hit against the int32 scalar oracle.  Returns the `parity` object."""
    from oracle import oracle as O
    tab, codes, off = w["tab"], w["codes"], w["off"]
    smax = int(np.asarray(tab).max())
    checked = mism = over_checked = 0
    vs = []
    if cpu is not None and gpu_scores:
        c, o, idx = cpu.sample
        for qi, ref in zip(w["cpu_queries"], cpu.last_scores):
            if qi not in gpu_scores:
                continue
            g = gpu_scores[qi][idx]
            r = ref.astype(np.int32)
            exact = r <= 32767 - smax if cpu.use_ref else np.ones(len(r), bool)
            mism += int((g[exact] != r[exact]).sum()); checked += int(exact.sum())
            for j in np.nonzero(~exact)[0][:256]:                 # the reference wrapped: int32 oracle decides
                t = c[o[j]:o[j + 1]]
                mism += int(int(g[j]) != O.score32(w["queries"][qi], t, tab, GAP_OPEN, GAP_EXTEND)); over_checked += 1
        vs.append("oracle/_ref libalign_ref.so on the cpu_baseline sample" if cpu.use_ref else "oracle/liboracle.so on the cpu_baseline sample")
    hits_checked = 0
    if hit_lists is not None and codes is not None:
        for q, hl in zip(w["queries"], hit_lists):
            if not len(hl):
                continue
            ent = np.array([e for _, e in hl], dtype=np.int64)
            lens = (off[1:] - off[:-1])[ent]
            o2 = np.zeros(len(ent) + 1, dtype=np.int64); np.cumsum(lens, out=o2[1:])
            c2 = np.concatenate([codes[off[e]:off[e + 1]] for e in ent])
            want = O.search32(q, c2, o2, tab, GAP_OPEN, GAP_EXTEND, threads=threads)
            got = np.array([s for s, _ in hl], dtype=np.int32)
            mism += int((got != want).sum()); hits_checked += len(ent)
        vs.append("oracle/liboracle.so score32 on every reported top hit")
    return {"checked": checked + over_checked + hits_checked, "mismatch": mism, "entries_vs_reference": checked,
            "entries_vs_int32_oracle": over_checked, "top_hits_vs_int32_oracle": hits_checked, "vs": "; ".join(vs) or None}


def run_reference_arm(a, w):
    """--impl reference: the reference's own CPU implementation of the path on the box's host cores."""
    cpu = CpuReference(w)
    per_step = max(1.0, min(a.cpu_seconds, 150.0 / max(a.steps + a.warmup, 1)))
    cpu.calibrate(per_step)
    secs = cells = 0.0
    for i in range(a.warmup + a.steps):
        s, c = cpu.run()
        if i >= a.warmup:
            secs += s; cells += c
    v = cells / max(secs, 1e-9) / 1e9
    one = cpu.one_thread()
    one.pop("seconds", None)
    line = {"metric": "GCUPS", "value": v, "unit": "GCUPS", "impl": "reference", "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": 1e3 * secs / max(a.steps, 1), "higher_is_better": True, "scaling": w["scaling"],
            "vs_baseline": None, "dtype": "int16", "data": "synthetic", "config": w["config"],
            "ran": {"what": "each step = one pass of the reference's kernel over a bounded sample of the workload above "
                            "(ms_per_step is the time of that pass, not of the whole workload)", "sample": cpu.describe(),
                    "host_cores_available": host_threads(), "omp_num_threads_env": os.environ.get("OMP_NUM_THREADS")},
            "cpu_baseline": {"value": v, "unit": "GCUPS", "kind": "reference" if cpu.use_ref else "port", "cores": cpu.threads,
                             "sample": cpu.describe(), "one_thread": one},
            "e2e": {"value": v, "unit": "GCUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line))
    return 0


# ----------------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scale", type=float, default=1.0, help="database size as a fraction of the named one (dev only; 1.0 = the benchmark)")
    ap.add_argument("--workload", default="c2", choices=["c1", "c2", "c3", "c4", "c5"],
                    help="c2 = the benchmark line; c1/c3/c4 = the other BASELINE configs; c5 = device-generated 10 G-residue database")
    ap.add_argument("--sharded", action="store_true", help="c2 only: ONE database of --shards-total x 205 M residues dealt over the ranks")
    ap.add_argument("--shards-total", type=int, default=8)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if a.sharded and a.workload != "c2":
        raise SystemExit("--sharded applies to --workload c2")

    if a.impl == "reference":
        if rank != 0:
            return 0
        if a.workload == "c5" or a.sharded:
            print(json.dumps({"impl": "reference", "unavailable": "this database exists on the device alone (generated there); the "
                              "reference arm is timed on the host-resident workloads c1-c4"}))
            return 0
        return run_reference_arm(a, load_workload(a, 0, 1))

    # ---------------- our arm ----------------
    import torch
    import swgpu
    if not torch.cuda.is_available() or swgpu.lib().swgpu_device_count() < 1:
        raise SystemExit("bench.py: no CUDA device -- the engine has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        # stdout carries the one JSON line and nothing else: NCCL prints its version banner there at NCCL_DEBUG=VERSION/INFO
        if os.environ.get("NCCL_DEBUG", "").upper() in ("VERSION", "INFO", "TRACE") and not os.environ.get("NCCL_DEBUG_FILE"):
            os.environ["NCCL_DEBUG"] = "WARN"
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    hbm_peak, sm_max, peak_src = load_peaks()
    threads_here = max(1, host_threads() // max(world, 1))

    t0 = time.time()
    w = load_workload(a, rank, world)
    queries, tab, device_db = w["queries"], w["tab"], w["codes"] is None
    t_gen = time.time() - t0
    t0 = time.time()
    if device_db:
        db = swgpu.Database.synthetic(w["lens"], w["seed"], entry_ids=w["entries"], device=local)
        n, n_res = len(w["lens"]), int(db.n_residues)
    else:
        db = swgpu.Database(w["codes"], w["off"], device=local, entry_base=w["entry_base"])
        n, n_res = len(w["off"]) - 1, int(w["off"][-1])
    t_pack = time.time() - t0
    cells_step = sum(len(q) for q in queries) * n_res
    nq = len(queries)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    from swgpu.shard import gather_topk_batch
    entries = w["entries"]
    base = 0 if device_db else w["entry_base"]

    def merge(hits):
        """per-rank top-k lists of the whole step -> ONE all_gather -> deterministic merge (score desc, entry asc)."""
        loc = np.full((len(hits), TOPK, 2), -1, dtype=np.int64)
        for qi, hl in enumerate(hits):
            for j, h in enumerate(hl):
                loc[qi, j, 0] = h["score"]
                loc[qi, j, 1] = int(entries[h["entry"]]) if entries is not None else base + h["entry"]
        return gather_topk_batch(loc, TOPK, dist, device=f"cuda:{local}" if dist is not None else None)

    last = {}

    def step_resident(qsub=None):
        """The queries of a step through swgpu_search_many(): up to three in flight, scores stay in HBM, top-k with
        coordinates per query; then ONE exchange + merge for the step (NCCL all_gather when N > 1)."""
        _, hits, per, bs = db.search_many(qsub or queries, tab, GAP_OPEN, GAP_EXTEND, k=TOPK, want_scores=False)
        merged, ex_ms, mg_ms = merge(hits)
        last["hits"], last["merged"] = hits, merged
        return bs["ms_total"] + ex_ms + mg_ms, bs["kernel_launches"], bs["n_rescored"], ex_ms + mg_ms

    # the serial per-kernel pass: all queries, or five spanning the lengths when there are 64 (c5)
    serial_queries = [queries[i] for i in (0, 5, 9, 11, 19)] if nq > 20 else queries

    def step_serial():
        """Same queries, one at a time (nothing overlaps): the pass the per-kernel K1S time is taken from."""
        ms, k1ms, k1cells, sms, rms = 0.0, 0.0, 0, 0.0, 0.0
        for q in serial_queries:
            st = db.search_resident(q, tab, GAP_OPEN, GAP_EXTEND)
            db.topk(TOPK)
            ms += st["ms_total"] + db.last_topk_ms
            k1ms += st["ms_inter"]; k1cells += st["cells_inter"]; sms += st["ms_total"]; rms += st["ms_rescore"]
        return ms, k1ms, k1cells, sms, rms

    want_e2e_scores = not device_db and n * nq * 4 <= (1 << 30)
    pinned = [torch.empty(n, dtype=torch.int32).pin_memory().numpy() for _ in queries] if want_e2e_scores else None

    def step_e2e():
        t = time.perf_counter()
        db.search_many(queries, tab, GAP_OPEN, GAP_EXTEND, k=0, out=pinned)
        return time.perf_counter() - t

    for _ in range(a.warmup):
        step_resident(queries[:3] if nq > 20 else None)
    barrier()
    sampler = ClockSampler(local); sampler.start(); time.sleep(0.25)
    tw0 = time.time()
    dev_ms, launches, resc, wall_steps, merge_ms = 0.0, 0, 0, 0.0, 0.0
    for _ in range(a.steps):
        t = time.perf_counter()
        r = step_resident()
        wall_steps += time.perf_counter() - t
        dev_ms += r[0]; launches += r[1]; resc += r[2]; merge_ms += r[3]
    barrier()
    tw1 = time.time()
    clocks = sampler.stop(tw0, tw1)
    hits_last, merged_last = last["hits"], last["merged"]

    # per-kernel time of the dominant kernel: the same queries, serially (in the pipelined region above the
    # event intervals of consecutive queries overlap, so they cannot be added up); a few steps are enough
    ser_steps = 1 if nq > 20 else min(a.steps, 5)
    ser_ms, k1ms, k1cells, search_ms, resc_ms = 0.0, 0.0, 0, 0.0, 0.0
    for _ in range(ser_steps):
        r = step_serial()
        ser_ms += r[0]; k1ms += r[1]; k1cells += r[2]; search_ms += r[3]; resc_ms += r[4]
    barrier()

    # e2e: host buffers in, host buffers out, wall clock.  Device-generated databases: 64 score vectors of 32.8 M entries
    # are not what a caller of a database this size reads back -- the e2e figure is the host wall clock of the timed steps
    # themselves (host queries and table in, top-100 hits with coordinates per query out to host memory, merge included).
    if want_e2e_scores:
        step_e2e()
        barrier()
        e2e_s = 0.0
        for _ in range(a.steps):
            e2e_s += step_e2e()
        barrier()
    else:
        e2e_s = wall_steps

    # CPU legs on rank 0 (single-GPU runs time the reference beside the GPU; every run checks parity)
    cpu = cpu_line = None
    have_cpu_db = not device_db and len(w["cpu_queries"]) > 0
    if have_cpu_db and rank == 0 and not (a.no_cpu and a.no_parity):
        cpu = CpuReference(w, threads=threads_here)
        cpu.calibrate(a.cpu_seconds if (world == 1 and not a.no_cpu) else min(a.cpu_seconds, 3.0))
        secs, cells = cpu.run()
        if world == 1 and not a.no_cpu:
            one = cpu.one_thread(); one.pop("seconds", None)
            cpu_line = {"value": cells / max(secs, 1e-9) / 1e9, "unit": "GCUPS", "cores": cpu.threads,
                        "kind": "reference" if cpu.use_ref else "port", "sample": cpu.describe(), "one_thread": one}
    parity = None
    if rank == 0 and not a.no_parity:
        if device_db:
            parity = {"checked": 0, "mismatch": 0, "vs": None,
                      "note": "database generated on the device: parity of this workload is the sampled oracle check of "
                              "tests/test_gpu_c5_device.py (bench.py holds no host copy of the residues)"}
        else:
            gpu_scores = {qi: pinned[qi] for qi in w["cpu_queries"]} if pinned is not None else {}
            parity = check_parity(w, cpu, gpu_scores, [[(h["score"], h["entry"]) for h in hl] for hl in hits_last], threads_here)

    if dist is not None:
        v = torch.tensor([dev_ms, e2e_s, tw1 - tw0, merge_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(v, op=dist.ReduceOp.MAX)
        dev_ms_max, e2e_max, wall_max, merge_ms_max = (float(x) for x in v)
        c = torch.tensor([cells_step], dtype=torch.float64, device="cuda"); dist.all_reduce(c)
        cells_all = float(c)
    else:
        dev_ms_max, e2e_max, wall_max, merge_ms_max, cells_all = dev_ms, e2e_s, tw1 - tw0, merge_ms, float(cells_step)

    rc = 0
    if rank == 0:
        gcups = cells_all * a.steps / (dev_ms_max * 1e-3) / 1e9
        alu_rate, alu_src = int_pipe_peak()
        sm_clk = (clocks.get("sm_mhz") or sm_max)
        # K1 cell roofline: 4.5 ALU-pipe (max-type DPX) warp-instructions per 64 cells (a warp-wide pair column),
        # i.e. 2.25 thread-instructions per cell, at the measured ALU-pipe rate and the max SM clock.
        sms = torch.cuda.get_device_properties(local).multi_processor_count
        peak_gcups = alu_rate * 32 * sms * sm_max * 1e6 / 2.25 / 1e9
        k1_gcups = k1cells / (k1ms * 1e-3) / 1e9 if k1ms > 0 else 0.0
        try:
            with open(os.path.join(ROOT, "profiles", "k1s_traffic.json")) as f:
                traffic = json.load(f)
        except OSError:
            traffic = {}
        # algorithmic HBM bytes of K1 per step: 1 B per target residue per query pass + 4 B score per target per query
        k1_bytes = ser_steps * len(serial_queries) * (n_res + 4 * n)
        hbm_gbs = k1_bytes / (k1ms * 1e-3) / 1e9 if k1ms > 0 else 0.0
        line = {
            "metric": "GCUPS", "value": gcups, "unit": "GCUPS", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": dev_ms_max / a.steps, "higher_is_better": True, "scaling": w["scaling"], "vs_baseline": None,
            "dtype": "u16x2 (packed 16-bit, 32-bit rescoring of saturated targets)", "data": "synthetic", "config": w["config"],
            "clocks": clocks, "gpu_launches": launches,
            "e2e": {"value": cells_all * a.steps / e2e_max / 1e9, "unit": "GCUPS",
                    "h2d_bytes_per_step": int(sum(len(q) + 1024 for q in queries)),
                    "d2h_bytes_per_step": int(nq * 4 * n) if want_e2e_scores else int(nq * TOPK * 32),
                    "note": ("swgpu_search_many(): host queries/table in, all per-entry int32 scores of every query out to pinned host "
                             "memory (k = 0: no top-k -- `value` times top-100 + coordinates with the scores left in HBM, so the two "
                             "time different tails), host wall clock; bytes are per rank; database packed and HBM-resident before the "
                             "timed region (one-off cost in setup_s; the reference does not time parsing either)") if want_e2e_scores else
                            "swgpu_search_many(): host queries/table in, top-100 hits with coordinates per query out to host memory, "
                            "merge included, host wall clock of the timed steps; bytes are per rank; database generated on the device "
                            "and HBM-resident (one-off cost in setup_s)"},
            "roofline": {"bound": "int-alu", "kernel": "k1s", "achieved": k1_gcups, "peak": peak_gcups, "unit": "GCUPS",
                         "frac": k1_gcups / peak_gcups if peak_gcups else None,
                         "peak_how": f"{alu_rate:.3f} warp-inst/clk/SM {alu_src} x 32 lanes x {sms} SMs x {sm_max:.0f} MHz / 2.25 ALU-pipe "
                                     f"thread-instructions per cell",
                         "sm_mhz_seen": sm_clk, "k1_ms_per_step": k1ms / ser_steps,
                         "k1_share_of_step": k1ms / ser_ms if ser_ms else None,
                         "k1_timing": "CUDA events around K1S in a serial pass of the same queries right after the timed region (one query "
                                      "at a time, nothing overlaps: k1_ms_per_step can therefore exceed the pipelined ms_per_step)",
                         "serial_ms_per_step": ser_ms / ser_steps, "serial_steps": ser_steps,
                         "traffic": traffic.get("single_stripe_lq1000", {}).get("bytes"),
                         "traffic_note": "DRAM bytes of ONE k1s launch over the C2 database (ncu --set full, profiles/k1s_traffic.json): "
                                         "single-stripe Lq=1000; multi-stripe Lq=2005 moves "
                                         f"{traffic.get('multi_stripe_lq2005', {}).get('bytes')} B; algorithmic "
                                         f"{N_RESIDUES + 4 * N_TARGETS} B per launch",
                         "hbm": {"achieved": hbm_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_gbs / hbm_peak, "peak_src": peak_src}},
            "search_only": {"value": (k1cells / (search_ms * 1e-3) / 1e9) if search_ms > 0 else None, "unit": "GCUPS",
                            "ms_per_step": search_ms / ser_steps, "rescore_ms_per_step": resc_ms / ser_steps,
                            "topk_ms_per_step": (ser_ms - search_ms) / ser_steps,
                            "what": "the serial pass split up: all per-entry scores final in HBM (16-bit pass + 32-bit rescoring), "
                                    "WITHOUT top-100 selection and the coordinate passes that `value` also times; this rank only"},
            "cpu_baseline": cpu_line, "parity": parity,
            "merge": {"ms_per_step": merge_ms_max / a.steps, "inside_value": True,
                      "how": "one exchange per step: [queries, 100, (score, global entry)] int64 -> H2D + all_gather_into_tensor + D2H "
                             "(CUDA events) + host lexsort (wall clock); N = 1: the host lexsort alone"},
            "topk_crc": int(zlib.crc32(np.ascontiguousarray(merged_last).tobytes())),
            "rescored_per_step": resc / a.steps,
            "setup_s": {"generate": t_gen, "pack_upload": t_pack}, "wall_s_timed": wall_max,
            "db": {"targets_per_gpu": n, "residues_per_gpu": n_res, "device_bytes": int(db.device_bytes)},
        }
        print(json.dumps(line))
        if parity is not None and parity["mismatch"]:
            print(f"bench.py: PARITY FAILURE: {parity['mismatch']} of {parity['checked']} checked scores differ", file=sys.stderr)
            rc = 3
    if dist is not None:
        dist.destroy_process_group()
    db.close()
    return rc


def _main_quiet_stdout():
    """stdout carries the ONE JSON line and nothing else: while the bench runs, file descriptor 1 points at stderr (NCCL
    prints its version banner to stdout at NCCL_DEBUG=VERSION, other libraries may chat there too); print() is routed
    to the saved descriptor."""
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(saved, "w", buffering=1)
    try:
        return main()
    finally:
        sys.stdout.flush()


if __name__ == "__main__":
    sys.exit(_main_quiet_stdout())
