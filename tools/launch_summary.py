#!/usr/bin/env python
"""Per-kernel totals and shares of an ncu launch list (--metrics gpu__time_duration.sum --csv --log-file ...).
Usage: python tools/launch_summary.py profiles/r02_launches.csv "<the profiled command>" > profiles/r02_launch_summary.txt"""
import collections
import csv
import re
import sys


def main():
    path = sys.argv[1]
    cmd = sys.argv[2] if len(sys.argv) > 2 else "?"
    hdr, rows = None, []
    for r in csv.reader(open(path)):
        if not r:
            continue
        if r[0] == "ID":
            hdr = r
            continue
        if hdr and len(r) == len(hdr):
            rows.append(dict(zip(hdr, r)))
    agg = collections.OrderedDict()
    for d in rows:
        name = re.sub(r"\(.*", "", d["Kernel Name"])
        v = float(d["Metric Value"].replace(",", ""))
        us = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}.get(d["Metric Unit"], 1.0) * v
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += us
    tot = sum(a[1] for a in agg.values())
    print("# per-kernel totals of %s: ncu --metrics gpu__time_duration.sum --clock-control none over" % path)
    print("#   %s" % cmd)
    print("# Cold-cache, serialised launches: the SHARES are what compares with the bench line of the same command, not the absolute times.")
    print("%-90s %8s %12s %12s %7s" % ("kernel", "launches", "total us", "us/launch", "share"))
    for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-90s %8d %12.1f %12.2f %6.1f%%" % (k[:90], n, us, us / n, 100 * us / tot))


if __name__ == "__main__":
    main()
