#!/usr/bin/env python
"""bench.py -- GCUPS of the one-to-many Smith-Waterman scoring path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--scale F] [--workload c2|c5]
  torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...        (N > 1: one rank per GPU, NCCL)

Workload (config 2 of BASELINE.json, SURVEY.md 8(d) "C2"): a Swiss-Prot-shaped synthetic database
(570 000 sequences, ~205 M residues, log-normal lengths, one 35 213-aa target; seed 43 + rank), BLOSUM62,
gap open 10 / extend 1, and the classic 20-query length sweep 144..5478 aa.  One "step" = the whole sweep
over the HBM-resident database = 41 752 x n_residues cell updates per GPU.  N > 1 is weak scaling: every
rank holds its own 205 M-residue shard (the database shards by residue count, no data-path collective);
after each step the per-rank top-k lists are merged with one small NCCL all_gather.

--workload c5 (config 5 of BASELINE.json, not the default bench line): a UniRef50-shaped database of 10 G residues
(32.8 M targets) GENERATED ON THE DEVICE (swgpu_db_create_synthetic), dealt longest-first round-robin over the ranks
(strong scaling: the total is fixed), 64 queries cycling through the same 20 lengths, BLOSUM45, top-100 per query merged
over NCCL.  One step = the 64 queries; a warm-up step = the first three.  --scale shrinks the database.

value  = GCUPS, device-timed (CUDA events on the engine's streams: first kernel launch of the sweep -> last
         top-k result of the sweep), real unpadded cells of all ranks / max-over-ranks time.  The sweep goes
         through swgpu_search_many(): up to three queries in flight, so one query's tail overlaps the next.
e2e    = same metric through the host-facing C-ABI call swgpu_search_many() with HOST buffers: queries + table
         in host memory, H2D inside, kernels, D2H of all per-entry scores of every query into pinned host
         memory, host wall clock.
roofline = dominant kernel (K1, packed 16-bit inter-sequence) against the measured integer-ALU-pipe
         cell-update roofline (DESIGN.md "Roofline"); the HBM view is reported beside it.
cpu_baseline = the reference's own AVX2/OpenMP kernel (oracle/_ref/libalign_ref.so, the unmodified
         reference sources) on the box's host cores, on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "seq-align-gpu_b200"))

GAP_OPEN, GAP_EXTEND = -10, -1
N_TARGETS, N_RESIDUES = 570_000, 205_000_000
TOPK = 100


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


def make_workload(rank, scale):
    from swgpu import synth
    n = max(64, int(N_TARGETS * scale))
    total = int(N_RESIDUES * scale)
    codes, off = synth.make_db(43 + rank, n, lo=2, hi=35213, total=total, force_max=35213 if scale >= 0.01 else None)
    queries = [synth.make_query(100 + i, L) for i, L in enumerate(synth.C2_QUERY_LENGTHS)]
    tab, _ = synth.load_matrix(os.path.join(ROOT, "scoring", "BLOSUM62.txt"))
    return codes, off, queries, tab


def make_workload_c5(rank, world, scale):
    """Lengths on the host, residues on the device.  Returns (lens of this rank's shard, global entry ids of the shard
    or None, queries, table, total residues of this shard)."""
    from swgpu import synth
    lens = synth.c5_lengths(scale)
    queries, _ = synth.c5_queries(lens, 64)
    tab, _ = synth.load_matrix(os.path.join(ROOT, "scoring", "BLOSUM45.txt"))
    entries = None
    if world > 1:
        order = np.argsort(-lens.astype(np.int64), kind="stable")
        entries = np.sort(order[rank::world])
        lens = lens[entries]
    return lens, entries, queries, tab


def cpu_sample(codes, off, stride):
    """Every stride-th GROUP of 16 consecutive targets of the length-sorted database: the reference batches 16
    consecutive records and pads them to the first one (src/alignment_cmdline.c:429-459), so sampling whole
    groups keeps its batch composition (and its padding cost) exactly what it is on the full database."""
    n = len(off) - 1
    groups = np.arange(stride // 2, n // 16, stride)        # mid-stride start: the one extreme batch (the 35 213-aa target,
                                                            # 16 lanes padded to it) is not over-weighted by a sparse sample
    idx = (groups[:, None] * 16 + np.arange(16)[None, :]).reshape(-1)
    lens = (off[1:] - off[:-1])[idx]
    o = np.zeros(len(idx) + 1, dtype=np.int64); np.cumsum(lens, out=o[1:])
    c = np.concatenate([codes[off[i]:off[i + 1]] for i in idx]) if len(idx) else np.zeros(0, np.int8)
    return c, o


def run_cpu_reference(codes, off, queries, tab, budget_s, qsel=(0, 3, 9)):
    """Times the reference's own kernel (unmodified sources, oracle/_ref) -- or, if that library did not
    travel, the C restatement -- with all host threads on a bounded sample.  Returns dict."""
    from oracle import oracle as O
    use_ref = O.have_ref()
    threads = O.ref().refh_max_threads() if use_ref else O.oracle().swo_max_threads()
    qs = [queries[i] for i in qsel]

    def once(c, o):
        secs, cells = 0.0, 0
        for q in qs:
            if use_ref:
                _, s = O.ref_search16(q, c, o, tab, GAP_OPEN, GAP_EXTEND, threads=threads)
            else:
                t0 = time.perf_counter(); O.search32(q, c, o, tab, GAP_OPEN, GAP_EXTEND, threads=threads); s = time.perf_counter() - t0
            secs += s; cells += len(q) * int(o[-1])
        return secs, cells

    stride = 512
    c, o = cpu_sample(codes, off, stride)
    s, cells = once(c, o)                                  # calibration (also warms the thread pool)
    rate = cells / max(s, 1e-6)
    want_cells = rate * budget_s
    full = sum(len(q) for q in qs) * int(off[-1])
    stride = int(max(1, min(512, round(full / max(want_cells, 1)))))
    c, o = cpu_sample(codes, off, stride)
    s, cells = once(c, o)
    return {"value": cells / s / 1e9, "unit": "GCUPS", "cores": int(threads), "kind": "reference" if use_ref else "port",
            "seconds": s,
            "sample": f"every {stride}-th batch of 16 consecutive targets (starting at batch {stride // 2}) of the length-sorted C2 database "
                      f"({len(o) - 1} targets, {int(o[-1])} residues) "
                      f"x queries of {', '.join(str(len(q)) for q in qs)} aa; timer = the reference's own region "
                      f"(omp parallel for over alignment_fill_matrices)"}, (c, o, qs)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scale", type=float, default=1.0, help="database size as a fraction of C2 (dev only; 1.0 = the benchmark)")
    ap.add_argument("--workload", default="c2", choices=["c2", "c5"], help="c2 = the benchmark line; c5 = device-generated 10 G-residue database")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu", action="store_true")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    config = {"workload": "C2: Swiss-Prot-shaped synthetic DB (570k seqs, ~205M residues per GPU), BLOSUM62, gap open 10 / extend 1, "
                          "20-query sweep 144..5478 aa; one step = the whole sweep",
              "targets_per_gpu": int(N_TARGETS * a.scale), "queries": 20, "topk": TOPK,
              "cache": "database + scratch per GPU (>=0.6 GB) exceed L2 (126 MB); every query pass streams the whole DB",
              "sharding": "one 205M-residue shard per rank, no data-path collective; NCCL all_gather of per-rank top-k"}

    c5 = a.workload == "c5"
    if c5:
        config = {"workload": f"C5: UniRef50-shaped synthetic DB generated on the device ({a.scale:g} x 10 G residues in total, "
                              "LogNormal(ln 230, 0.75) lengths), BLOSUM45, gap open 10 / extend 1, 64 queries cycling through "
                              "144..5478 aa, top-100 per query; one step = the 64 queries",
                  "queries": 64, "topk": TOPK,
                  "cache": "the database per GPU exceeds L2 (126 MB) by orders of magnitude; every query pass streams it",
                  "sharding": "targets dealt longest-first round-robin over the ranks (total fixed: strong scaling), no data-path "
                              "collective; NCCL all_gather of per-rank top-k"}
    if a.impl == "reference":
        if rank != 0:
            return 0
        if c5:
            print(json.dumps({"impl": "reference", "unavailable": "the reference arm is timed on the C2 workload only "
                              "(C5 exists on the device alone; a 1/1024 CPU extrapolation is not a measurement)"}))
            return 0
        codes, off, queries, tab = make_workload(0, a.scale)
        # each step is a bounded sample sized so steps+warmup finish within a few minutes
        per_step = max(2.0, min(a.cpu_seconds, 150.0 / max(a.steps + a.warmup, 1)))
        vals, info = [], None
        for i in range(a.warmup + a.steps):
            info, _ = run_cpu_reference(codes, off, queries, tab, per_step)
            if i >= a.warmup:
                vals.append(info)
        cells = sum(v["value"] * v["seconds"] for v in vals); secs = sum(v["seconds"] for v in vals)
        v = cells / secs
        line = {"metric": "GCUPS", "value": v, "unit": "GCUPS", "impl": "reference", "n_gpus": a.gpus, "steps": a.steps,
                "warmup": a.warmup, "ms_per_step": 1e3 * secs / max(a.steps, 1), "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "int16", "data": "synthetic", "config": config,
                "cpu_baseline": {k: info[k] for k in ("kind", "cores", "sample")} | {"value": v, "unit": "GCUPS"},
                "e2e": {"value": v, "unit": "GCUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
        print(json.dumps(line)); return 0

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

    t0 = time.time()
    entries = None
    if c5:
        from swgpu import synth
        lens, entries, queries, tab = make_workload_c5(rank, world, a.scale)
        t_gen = time.time() - t0
        t0 = time.time()
        db = swgpu.Database.synthetic(lens, synth.C5_SEED, entry_ids=entries, device=local)
        t_pack = time.time() - t0
        n, n_res = len(lens), int(db.n_residues)
    else:
        codes, off, queries, tab = make_workload(rank, a.scale)
        t_gen = time.time() - t0
        t0 = time.time()
        db = swgpu.Database(codes, off, device=local, entry_base=rank * (len(off) - 1))
        t_pack = time.time() - t0
        n = len(off) - 1
        n_res = int(off[-1])
    cells_step = sum(len(q) for q in queries) * n_res

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    from swgpu.shard import gather_topk

    def merge_topk(hits_local):
        """per-rank top-k -> NCCL all_gather -> deterministic merge (score desc, entry asc)."""
        gid = (lambda e: int(entries[e])) if entries is not None else (lambda e: e)
        return gather_topk([[h["score"], gid(h["entry"])] for h in hits_local], TOPK, dist, device="cuda" if dist is not None else None)

    def step_resident(qsub=None):
        """The sweep through swgpu_search_many(): up to three queries in flight, scores stay in HBM, top-k with
        coordinates per query; then the per-query merge (NCCL all_gather when N > 1)."""
        _, hits, per, bs = db.search_many(qsub or queries, tab, GAP_OPEN, GAP_EXTEND, k=TOPK, want_scores=False)
        for h in hits:
            merge_topk(h)
        return bs["ms_total"], bs["kernel_launches"], bs["n_rescored"]

    # C5: the serial per-kernel pass runs five queries spanning the lengths (the whole 64 would double the run)
    serial_queries = [queries[i] for i in (0, 5, 9, 11, 19)] if c5 else queries

    def step_serial():
        """Same sweep, one query at a time (nothing overlaps): the pass the per-kernel K1S time is taken from."""
        ms, k1ms, k1cells = 0.0, 0.0, 0
        for q in serial_queries:
            st = db.search_resident(q, tab, GAP_OPEN, GAP_EXTEND)
            db.topk(TOPK)
            ms += st["ms_total"] + db.last_topk_ms
            k1ms += st["ms_inter"]; k1cells += st["cells_inter"]
        return ms, k1ms, k1cells

    pinned = None if c5 else [torch.empty(n, dtype=torch.int32).pin_memory().numpy() for _ in queries]

    def step_e2e():
        t = time.perf_counter()
        db.search_many(queries, tab, GAP_OPEN, GAP_EXTEND, k=0, out=pinned)
        return time.perf_counter() - t

    for _ in range(a.warmup):
        step_resident(queries[:3] if c5 else None)
    barrier()
    sampler = ClockSampler(local); sampler.start(); time.sleep(0.25)
    tw0 = time.time()
    dev_ms, launches, resc, wall_steps = 0.0, 0, 0, 0.0
    for _ in range(a.steps):
        t = time.perf_counter()
        r = step_resident()
        wall_steps += time.perf_counter() - t
        dev_ms += r[0]; launches += r[1]; resc += r[2]
    barrier()
    tw1 = time.time()
    clocks = sampler.stop(tw0, tw1)

    # per-kernel time of the dominant kernel: the same K steps, serially (in the pipelined region above the
    # event intervals of consecutive queries overlap, so they cannot be added up)
    ser_ms, k1ms, k1cells = 0.0, 0.0, 0
    for _ in range(1 if c5 else a.steps):
        r = step_serial()
        ser_ms += r[0]; k1ms += r[1]; k1cells += r[2]
    barrier()

    # e2e: host buffers in, host buffers out, wall clock.  C5: 64 score vectors of 32.8 M entries are not what a caller of a
    # database this size reads back -- the e2e figure is the host wall clock of the timed steps themselves (host queries and
    # table in, top-100 hits with coordinates per query out to host memory, merge included).
    if c5:
        e2e_s = wall_steps
    else:
        step_e2e()
        barrier()
        e2e_s = 0.0
        for _ in range(a.steps):
            e2e_s += step_e2e()
        barrier()

    if dist is not None:
        v = torch.tensor([dev_ms, e2e_s, tw1 - tw0], dtype=torch.float64, device="cuda")
        dist.all_reduce(v, op=dist.ReduceOp.MAX)
        dev_ms_max, e2e_max, wall_max = (float(x) for x in v)
        c = torch.tensor([cells_step], dtype=torch.float64, device="cuda"); dist.all_reduce(c)
        cells_all = float(c)
    else:
        dev_ms_max, e2e_max, wall_max, cells_all = dev_ms, e2e_s, tw1 - tw0, float(cells_step)

    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu and not c5:
        cpu, _ = run_cpu_reference(codes, off, queries, tab, a.cpu_seconds)
        cpu.pop("seconds", None)

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
        k1_bytes = (1 if c5 else a.steps) * len(serial_queries) * (n_res + 4 * n)
        hbm_gbs = k1_bytes / (k1ms * 1e-3) / 1e9 if k1ms > 0 else 0.0
        line = {
            "metric": "GCUPS", "value": gcups, "unit": "GCUPS", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": dev_ms_max / a.steps, "higher_is_better": True, "scaling": "strong" if c5 else "weak", "vs_baseline": None,
            "dtype": "u16x2 (packed 16-bit, 32-bit rescoring of saturated targets)", "data": "synthetic", "config": config,
            "clocks": clocks, "gpu_launches": launches,
            "e2e": {"value": cells_all * a.steps / e2e_max / 1e9, "unit": "GCUPS",
                    "h2d_bytes_per_step": int(sum(len(q) + 1024 for q in queries)),
                    "d2h_bytes_per_step": int(len(queries) * TOPK * 32) if c5 else int(len(queries) * 4 * n),
                    "note": ("swgpu_search_many(): host queries/table in, top-100 hits with coordinates per query out to host memory, "
                             "merge included, host wall clock of the timed steps; bytes are per rank; database generated on the device "
                             "and HBM-resident (one-off cost in setup_s)") if c5 else
                            "swgpu_search_many(): host queries/table in, all per-entry int32 scores of every query out to pinned host "
                            "memory, host wall clock; bytes are per rank; database packed and HBM-resident (one-off cost in setup_s)"},
            "roofline": {"bound": "int-alu", "kernel": "k1s", "achieved": k1_gcups, "peak": peak_gcups, "unit": "GCUPS",
                         "frac": k1_gcups / peak_gcups if peak_gcups else None,
                         "peak_how": f"{alu_rate:.3f} warp-inst/clk/SM {alu_src} x 32 lanes x {sms} SMs x {sm_max:.0f} MHz / 2.25 ALU-pipe "
                                     f"thread-instructions per cell",
                         "sm_mhz_seen": sm_clk, "k1_ms_per_step": k1ms / (1 if c5 else a.steps),
                         "k1_share_of_step": k1ms / ser_ms if ser_ms else None,
                         "k1_timing": "CUDA events around K1S in a serial pass of the same K steps right after the timed region "
                                      "(one query at a time; in the pipelined region consecutive queries overlap)",
                         "serial_ms_per_step": ser_ms / (1 if c5 else a.steps),
                         "traffic": traffic.get("single_stripe_lq1000", {}).get("bytes"),
                         "traffic_note": "DRAM bytes of ONE k1s launch over this database (ncu --set full, profiles/k1s_traffic.json): "
                                         "single-stripe Lq=1000; multi-stripe Lq=2005 moves "
                                         f"{traffic.get('multi_stripe_lq2005', {}).get('bytes')} B; algorithmic "
                                         f"{n_res + 4 * n} B per launch",
                         "hbm": {"achieved": hbm_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_gbs / hbm_peak, "peak_src": peak_src}},
            "cpu_baseline": cpu,
            "rescored_per_step": resc / a.steps,
            "setup_s": {"generate": t_gen, "pack_upload": t_pack}, "wall_s_timed": wall_max,
            "db": {"targets_per_gpu": n, "residues_per_gpu": n_res, "device_bytes": int(db.device_bytes)},
        }
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()
    db.close()
    return 0


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
