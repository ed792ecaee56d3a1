"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import collections
import csv
import sys


def main(path, note=""):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows[1:]:
        v = float(r[vi].replace(",", ""))
        v = v / 1000 if r[ui] == "ns" else (v * 1000 if r[ui] == "ms" else v)
        agg[r[ki][:110]][0] += 1
        agg[r[ki][:110]][1] += v
    tot = sum(v[1] for v in agg.values())
    print(f"# {path}: {len(rows) - 1} launches. {note}")
    print("# per-launch times under ncu are cold-cache and serialised: compare SHARES, not absolutes")
    for k, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"{t:9.1f}us {100 * t / tot:5.1f}% n={c:4d} avg={t / c:7.2f}us  {k}")
    print(f"total {tot:.1f} us")


if __name__ == "__main__":
    main(sys.argv[1], " ".join(sys.argv[2:]))
