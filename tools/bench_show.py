#!/usr/bin/env python
"""Print the per-kernel table of bench.py JSON lines: python tools/bench_show.py file.json [...]"""
import json
import sys

for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable:", e)
        continue
    print(f"{f}: {d['value']:.2f} {d['unit']}, {d['ms_per_step']:.2f} ms/step, e2e {d.get('e2e', {}).get('value')}, "
          f"launches {d.get('gpu_launches')}")
    print("   stages", {k: round(v, 2) for k, v in d["stages_ms"].items()})
    for k, v in sorted(d.get("kernels", {}).items(), key=lambda x: -x[1]["ms_per_step"]):
        extra = f"{v['achieved_gbs']:.0f} GB/s frac {v['frac']:.3f}" if "frac" in v else ""
        print(f"   {k:28s} {v['ms_per_step']:8.3f} ms {v['launches_per_step']:6.1f} launches  {extra}")
        for c, cv in v.get("classes", {}).items():
            extra = f"{cv['alg_bytes_per_step'] / 1e6:.1f} MB alg, {cv['achieved_gbs']:.0f} GB/s" if "achieved_gbs" in cv else ""
            print(f"        {c:26s} {cv['ms_per_step']:8.3f} ms {cv['launches_per_step']:6.1f}  {extra}")
