#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: total time and launches per kernel."""
import collections
import csv
import re
import sys


def main(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    tot = collections.defaultdict(float)
    cnt = collections.Counter()
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", row["Kernel Name"])
        v = float(row["Metric Value"].replace(",", ""))
        unit = row["Metric Unit"]
        v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(unit, 1e-6)
        tot[name] += v
        cnt[name] += 1
    total = sum(tot.values())
    print(f"{sum(cnt.values())} launches, {total:.3f} ms in kernels")
    print(f"{'ms':>10} {'share':>7} {'launches':>8}  kernel")
    for k, v in sorted(tot.items(), key=lambda x: -x[1]):
        print(f"{v:10.3f} {100 * v / total:6.1f}% {cnt[k]:8d}  {k}")


if __name__ == "__main__":
    main(sys.argv[1])
