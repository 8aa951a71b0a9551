#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per (kernel, grid) launches, mean / min duration in us.
usage: python tools/launch_summary.py launches.csv"""
import collections
import csv
import sys


def main():
    lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        try:
            v = float(row["Metric Value"].replace(",", ""))
        except (KeyError, ValueError):
            continue
        if row.get("Metric Unit", "ns").startswith("us"):
            v *= 1e3
        agg.setdefault((row["Kernel Name"][:48], row.get("Grid Size", "")), []).append(v)
    print("%-50s %-16s %5s %10s %10s %12s" % ("kernel", "grid", "n", "mean us", "min us", "total us"))
    for (k, g), v in agg.items():
        print("%-50s %-16s %5d %10.1f %10.1f %12.1f" % (k, g, len(v), sum(v) / len(v) / 1e3, min(v) / 1e3, sum(v) / 1e3))


if __name__ == "__main__":
    main()
