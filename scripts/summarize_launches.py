#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel.
usage: summarize_launches.py launches.csv ["command line that was profiled"]"""
import collections
import csv
import sys


def main():
    path = sys.argv[1]
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    scale = {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "s": 1e6, "second": 1e6, "nsecond": 1e-3}
    for r in rows[1:]:
        v = float(r[vi].replace(",", "")) * scale.get(r[ui], 1.0)
        a = agg.setdefault(r[ki], [0, 0.0])
        a[0] += 1
        a[1] += v
    total = sum(t for _, t in agg.values())
    if len(sys.argv) > 2:
        print("ncu --metrics gpu__time_duration.sum --clock-control none, %s" % sys.argv[2])
    print("per-launch times are cold-cache and serialised: compare shares, not absolutes\n")
    print("%-100s %6s %12s %10s %7s" % ("kernel", "count", "total_us", "mean_us", "share"))
    for k, (c, t) in agg.items():
        print("%-100s %6d %12.1f %10.1f %6.1f%%" % (k[:100], c, t, t / c, 100.0 * t / total))


if __name__ == "__main__":
    main()
