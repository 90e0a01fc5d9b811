#!/usr/bin/env python3
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel totals and one step's sequence."""
import collections
import csv
import re
import sys


def load(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    rows = list(csv.DictReader(lines))
    out = []
    for r in rows:
        v = float(r["Metric Value"].replace(",", ""))
        u = r["Metric Unit"]
        v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v * 1e6 if u == "s" else v
        name = re.sub(r"\(.*", "", r["Kernel Name"])
        name = name.replace("b200::<unnamed>::", "").replace("void ", "")
        out.append((name, r["Grid Size"], r["Block Size"], v))
    return out


def main():
    rows = load(sys.argv[1])
    only = len(sys.argv) > 2 and sys.argv[2] == "--ours"
    agg = collections.OrderedDict()
    for name, g, b, v in rows:
        if only and name.startswith("at::"):
            continue
        a = agg.setdefault(name[:64], [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print("%-66s %5s %12s %10s %6s" % ("kernel", "n", "total us", "avg us", "share"))
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-66s %5d %12.1f %10.1f %5.1f%%" % (k, n, t, t / n, 100 * t / tot))
    idx = [i for i, r in enumerate(rows) if "fusedXlateDecimKernel" in r[0] or "fusedStreamKernel" in r[0]]
    if len(idx) >= 3:
        s, e = idx[2], idx[3] if len(idx) > 3 else len(rows)
        print("\none step (launch order):")
        for name, g, b, v in rows[s - 1:e - 1]:
            print("  %-58s grid %-16s block %-12s %9.1f us" % (name[:58], g, b, v))


if __name__ == "__main__":
    main()
