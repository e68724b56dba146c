#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel launches, total time, share."""
import collections
import csv
import sys


def main(path):
    hdr = None
    agg = collections.defaultdict(lambda: [0, 0.0])
    unit = "ns"
    for r in csv.reader(open(path)):
        if "Kernel Name" in r:
            hdr = r
            continue
        if hdr and len(r) == len(hdr):
            d = dict(zip(hdr, r))
            if d["Metric Name"] == "gpu__time_duration.sum":
                name = d["Kernel Name"].split("(")[0]
                agg[name][0] += 1
                agg[name][1] += float(d["Metric Value"].replace(",", ""))
                unit = d["Metric Unit"]
    tot = sum(v[1] for v in agg.values())
    print(f"{'kernel':40s} {'launches':>8s} {'total_' + unit:>14s} {'mean_' + unit:>12s} {'share':>7s}")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:40s} {v[0]:8d} {v[1]:14.0f} {v[1] / v[0]:12.0f} {v[1] / tot:7.3f}")


if __name__ == "__main__":
    main(sys.argv[1])
