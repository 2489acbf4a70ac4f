#!/usr/bin/env python
"""one FM-index build (device-resident text) for ncu launch lists:  python tools/fm_profile_run.py [n] [reps]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import kiss_b200  # noqa: E402
import textgen as tg  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 400_000_000
S = tg.uniform_big(5, n)
ctx = kiss_b200.Context(0)
for rep in range(int(sys.argv[2]) if len(sys.argv) > 2 else 3):
    t0 = time.perf_counter()
    img = ctx.fmindex_build(S, 32)
    dt = time.perf_counter() - t0
    t = ctx.last_timings
    print(rep, "host call", round(dt, 4), "sa device", round(t["total_device"], 4), "fmindex", round(t["fmindex"], 4), flush=True)
