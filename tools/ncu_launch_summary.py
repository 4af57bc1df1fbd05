#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: time and share per kernel."""
import collections
import csv
import re
import sys


def short(name):
    m = re.search(r"(\w+_kernel\w*|\w+Kernel\w*|\w+)\s*(<[^(]*>)?\s*\(", name)
    base = m.group(1) if m else name
    tpl = re.search(r"<([^()]*)>\s*\(", name)
    if "frn::" in name and tpl:
        base += "<" + tpl.group(1).replace("(int)", "") + ">"
    return base[:80]


def main(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        us = v / 1e3 if u in ("ns", "nsecond") else v * 1e3 if u in ("ms", "msecond") else v
        k = short(row["Kernel Name"])
        agg[k][0] += 1
        agg[k][1] += us
    tot = sum(v[1] for v in agg.values())
    print(f"{'kernel':80s} {'launches':>8s} {'ms':>10s} {'share':>7s}")
    for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"{k:80s} {v[0]:8d} {v[1] / 1e3:10.3f} {100 * v[1] / tot:6.2f}%")
    print(f"{'TOTAL':80s} {sum(v[0] for v in agg.values()):8d} {tot / 1e3:10.3f}")


if __name__ == "__main__":
    main(sys.argv[1])
