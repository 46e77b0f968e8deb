#!/usr/bin/env python
"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list into per-kernel count / total / mean."""
import csv
import re
import sys
from collections import OrderedDict


def main():
    rows = []
    with open(sys.argv[1]) as f:
        lines = [l for l in f if l.startswith('"')]
    rd = csv.reader(lines)
    hdr = next(rd)
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    mi = hdr.index("Metric Name")
    agg = OrderedDict()
    for r in rd:
        if r[mi] != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", r[ki]).replace("void ", "").replace("<unnamed>::", "")
        name = re.sub(r"cub::CUB_\d+_\w+?::", "cub::", name)[:90]
        v = float(r[vi].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r[ui], 1.0)
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print(f"{'kernel':90s} {'launches':>8s} {'total us':>12s} {'mean us':>10s} {'share':>7s}")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:90s} {n:8d} {t:12.1f} {t / n:10.1f} {100 * t / tot:6.1f}%")
    print(f"{'total':90s} {sum(a[0] for a in agg.values()):8d} {tot:12.1f}")


if __name__ == "__main__":
    main()
