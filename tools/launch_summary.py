#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total and share.

  python tools/launch_summary.py gpurun_out/launches.csv [--skip N] [--last M]
"""
import csv
import re
import sys
from collections import OrderedDict


def main():
    path = sys.argv[1]
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if not l.startswith("==")]
    rd = csv.DictReader(lines)
    for r in rd:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        if unit in ("us", "usecond"):
            v *= 1e3
        elif unit in ("ms", "msecond"):
            v *= 1e6
        name = r["Kernel Name"]
        rows.append((int(r["ID"]), name, v, r.get("Grid Size", ""), r.get("Block Size", "")))
    last = None
    if "--last" in sys.argv:
        last = int(sys.argv[sys.argv.index("--last") + 1])
        rows = rows[-last:]
    agg = OrderedDict()
    for _, name, v, grid, blk in rows:
        short = re.sub(r"\(.*", "", name)
        key = short
        a = agg.setdefault(key, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print(f"{len(rows)} launches, total {tot/1e3:.1f} us")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{t/1e3:10.1f} us {100*t/tot:5.1f}%  n={n:4d}  avg {t/n/1e3:8.1f} us  {k}")


if __name__ == "__main__":
    main()
