import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, kiss_b200, textgen as tg
n = int(sys.argv[1]) if len(sys.argv) > 1 else 400_000_000
S = tg.chm13_like(3, n, base=tg.uniform_big(3, n))
ctx = kiss_b200.Context(0)
sa = ctx.suffix_array_dna(S, 256)
print(ctx.last_timings)
