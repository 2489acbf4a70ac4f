import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, kiss_b200, textgen as tg
N = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000_000
S = tg.chm13_like(3, N, base=tg.uniform_big(3, N))
ctx = kiss_b200.Context(0)
for rep in range(int(sys.argv[2]) if len(sys.argv) > 2 else 3):
    t0 = time.perf_counter(); sa = ctx.suffix_array_dna(S, 256, kiss_b200.PREFIX_DOUBLING); dt = time.perf_counter() - t0
    print(rep, "host call", round(dt, 4), "lms_sort", round(ctx.last_timings["lms_sort"], 4), "rounds", ctx.last_timings["lms_sort_rounds"], flush=True)
