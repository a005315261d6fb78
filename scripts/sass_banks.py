"""Per kernel of a cubin / executable: how many DPX instructions (VIMNMX3 / VIADDMNMX .U16x2) have 1, 2 or 3 distinct source
registers in the same register-file bank (even / odd register number).  Read here, no GPU.
  python scripts/sass_banks.py bench_micro/ubench_bank [name-filter]"""
import collections, re, subprocess, sys

txt = subprocess.run(["cuobjdump", "-sass", sys.argv[1]], capture_output=True, text=True).stdout
flt = sys.argv[2] if len(sys.argv) > 2 else ""
cur, res = None, {}
for l in txt.splitlines():
    m = re.search(r"Function : (\S+)", l)
    if m:
        cur = m.group(1); res[cur] = collections.Counter(); continue
    m = re.match(r"\s+/\*([0-9a-f]+)\*/\s+(@!?U?P\d\s+)?(\S+)\s+(.*);", l)
    if not m or cur is None:
        continue
    op = m.group(3)
    if not (op.startswith("VIADDMNMX.U16") or op.startswith("VIMNMX3.U16")):
        continue
    regs = set()
    for s in m.group(4).split(",")[1:]:
        mm = re.match(r"\s*-?R(\d+)", s)
        if mm:
            regs.add(int(mm.group(1)))
    ev = len([r for r in regs if r % 2 == 0])
    res[cur][(op.split(".")[0], f"{len(regs)} regs", f"max {max(ev, len(regs) - ev)} per bank")] += 1
for k, v in res.items():
    if flt in k and v:
        tot = sum(v.values()); bad = sum(n for (o, r, b), n in v.items() if b == "max 3 per bank")
        print(f"{k}: {tot} DPX, {bad} with three sources in one bank ({100.0 * bad / tot:.1f} %)")
