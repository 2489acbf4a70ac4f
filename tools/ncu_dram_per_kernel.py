#!/usr/bin/env python
"""Per kernel function: launches, total time, DRAM bytes read+written (sum and mean per launch) from
`ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv` output.
python tools/ncu_dram_per_kernel.py launches.csv [profiles/ncu_dram_per_launch.json workload]"""
import collections
import csv
import re
import sys


def main(path, json_out=None, workload=None):
    lines = [l for l in open(path) if not l.startswith("==")]
    t = collections.defaultdict(float)
    rd = collections.defaultdict(float)
    wr = collections.defaultdict(float)
    ids = collections.defaultdict(set)
    scale_t = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}
    scale_b = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    for row in csv.DictReader(lines):
        name = re.sub(r"\(.*", "", row["Kernel Name"])
        name = re.sub(r"<.*", "", name).replace("void ", "").replace("kiss::", "")
        v = float(row["Metric Value"].replace(",", ""))
        unit = row["Metric Unit"]
        ids[name].add(row["ID"])
        mn = row["Metric Name"]
        if mn == "gpu__time_duration.sum":
            t[name] += v * scale_t.get(unit, 1e-6)
        elif mn == "dram__bytes_read.sum":
            rd[name] += v * scale_b.get(unit, 1.0)
        elif mn == "dram__bytes_write.sum":
            wr[name] += v * scale_b.get(unit, 1.0)
    tot = sum(t.values())
    print(f"{sum(len(v) for v in ids.values())} launches, {tot:.3f} ms in kernels (serialised, cold cache: shares, not absolutes)")
    print(f"{'ms':>9} {'share':>6} {'launches':>8} {'dram GB':>9} {'GB/launch':>10}  kernel")
    for k, v in sorted(t.items(), key=lambda x: -x[1]):
        n = len(ids[k])
        b = rd[k] + wr[k]
        print(f"{v:9.3f} {100 * v / tot:5.1f}% {n:8d} {b / 1e9:9.3f} {b / n / 1e9:10.4f}  {k}")
    if json_out:  # per-launch DRAM bytes per kernel, keyed by workload: bench.py's `roofline.traffic` reads this file
        import json
        import os
        cur = json.load(open(json_out)) if os.path.exists(json_out) else {}
        cur[workload or "c3"] = {"source": os.path.basename(path),
                                 "dram_bytes_per_launch": {k: (rd[k] + wr[k]) / len(ids[k]) for k in t},
                                 "launches": {k: len(ids[k]) for k in t}}
        json.dump(cur, open(json_out, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main(*sys.argv[1:4])
