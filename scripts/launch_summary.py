"""Per-kernel totals of an ncu launch list (--metrics gpu__time_duration.sum --csv --log-file X.csv).
  python scripts/launch_summary.py profiles/r1d_launches_bench.csv"""
import collections, csv, re, sys
rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 10]
hdr = rows[0]; ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
n = collections.Counter(); ms = collections.Counter()
for r in rows[1:]:
    name = r[ik]
    name = re.sub(r"^void |\(.*$|swk::", "", name)
    if name.startswith("k1s<"):
        name = "k1s<R, lanes, multi> (all instantiations)"
    v = float(r[iv].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[iu], 1e-6)
    n[name] += 1; ms[name] += v
tot = sum(ms.values())
print(f"{sum(n.values())} launches, {tot:.1f} ms of kernel time (serialised, cold caches)")
print(f"{'kernel':<48}{'launches':>9}{'ms':>11}{'share':>8}")
for k, v in ms.most_common():
    print(f"{k:<48}{n[k]:>9}{v:>11.3f}{100 * v / tot:>7.2f}%")
