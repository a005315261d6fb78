"""Summarise an .ncu-rep (read here, no GPU): key raw metrics + per-opcode instruction counts and stall samples.
  python scripts/ncu_summary.py gpurun_out/x.ncu-rep [cells]"""
import collections, csv, io, re, subprocess, sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
d = {h: (u, v) for h, u, v in zip(hdr, units, vals)}
want = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.per_cycle_active", "smsp__warps_eligible.avg.per_cycle_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__bytes.sum.per_second", "lts__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "sm__cycles_elapsed.avg", "sm__cycles_active.avg"]
for k in want:
    if k in d:
        print(f"{k} | {d[k][0]} | {d[k][1]}")
for h, (u, v) in d.items():
    if re.search(r"smsp__average_warps_issue_stalled_.*_per_issue_active.ratio", h) and float(v or 0) > 0.02:
        print(f"{h.replace('smsp__average_warps_issue_stalled_', 'stall ').replace('_per_issue_active.ratio', '')} | {float(v):.3f}")

src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}
ops = collections.Counter(); samp = collections.Counter(); stall = collections.defaultdict(collections.Counter)
stallcols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = 0
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    s = r[ix["Source"]].split()
    op = (s[1] if s[0].startswith("@") else s[0]).split(".")[0]
    n = int(r[ix["Instructions Executed"]]); ops[op] += n; tot += n
    samp[op] += int(r[ix["# Samples"]])
    for c in stallcols:
        stall[op][c] += int(r[ix[c]])
ts = sum(samp.values())
cells = float(sys.argv[2]) if len(sys.argv) > 2 else None
print(f"total warp-inst {tot:.4g}" + (f" = {tot / (cells / 64):.2f} per warp-cellpair" if cells else ""))
for op, n in ops.most_common(16):
    top = ", ".join(f"{k[6:]}:{100 * v / max(samp[op], 1):.0f}%" for k, v in stall[op].most_common(3))
    per = f" {n / (cells / 64):5.2f}/wcp" if cells else ""
    print(f"  {op:10s} {100 * n / tot:5.1f}%{per}  samples {100 * samp[op] / ts:5.1f}%  {top}")
agg = collections.Counter()
for op in stall:
    agg.update(stall[op])
print("stall samples:", ", ".join(f"{k[6:]} {100 * v / ts:.1f}%" for k, v in agg.most_common(8)))
lines = sorted((r for r in rows[2:] if len(r) >= len(hdr)), key=lambda r: -int(r[ix["# Samples"]]))[:12]
for r in lines:
    print(f"  {int(r[ix['# Samples']]):7d} {r[ix['Source']].strip()[:64]:64s}", {c[6:]: int(r[ix[c]]) for c in stallcols if int(r[ix[c]]) > 0.2 * int(r[ix['# Samples']])})
