"""Per-source-line stall samples from `ncu -i X.ncu-rep --page source --csv --print-source sass,cuda -k <kernel>`.
usage: python tools/ncu_lines.py /tmp/trunk_sc.csv [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
names = None
data = []
files = []
fname = ""
for r in rows:
    if r and r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if r and r[0] == "Line No":
        names = r
        continue
    if names is None or len(r) < len(names) or r[2] != "-":   # keep the per-line aggregate rows (Address == "-")
        continue
    data.append([fname] + r[1:] if False else r)
    files.append(fname)
si = names.index("# Samples"); ie = names.index("Instructions Executed")
stall = [j for j, n in enumerate(names) if n.startswith("stall_") and "Not Issued" not in n]
xs = names.index("L1 Wavefronts Shared Excessive") if "L1 Wavefronts Shared Excessive" in names else si
tot = sum(int(r[si] or 0) for r in data)
print("total samples", tot)
for r, fn in sorted(zip(data, files), key=lambda x: -int(x[0][si] or 0))[:top]:
    s = int(r[si] or 0)
    st = sorted(((int(r[j] or 0), names[j][6:]) for j in stall), reverse=True)[:3]
    st = " ".join(f"{n}:{v}" for v, n in st if v)
    print(f"{s:6d} {100*s/max(tot,1):5.1f}% {fn[:16]:16s} L{r[0]:>4s} exec={r[ie]:>8s} xs={r[xs]:>7s} {r[1].strip()[:80]:80s} {st}")
