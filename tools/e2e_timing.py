#!/usr/bin/env python
"""Stage timings of the host API with pinned buffers (the overlapped copy-out path) on uniform text."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import kiss_b200  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000_000
    ctx = kiss_b200.Context(0)
    d = torch.empty(n + 64, dtype=torch.uint8, device="cuda")
    ctx.synth_uniform_dev(d.data_ptr(), n, 2)
    hS = torch.empty(n, dtype=torch.uint8).pin_memory()
    hS.copy_(d[:n])
    del d
    torch.cuda.empty_cache()
    hSA = torch.empty(n + 1, dtype=torch.int32).pin_memory()
    pageable = np.empty(n + 1, np.uint32)
    for name, out in (("pinned", hSA.numpy().view(np.uint32)), ("pinned", hSA.numpy().view(np.uint32)),
                      ("pageable", pageable)):
        t0 = time.perf_counter()
        ctx.suffix_array_dna(hS.numpy(), 256, out=out)
        wall = time.perf_counter() - t0
        t = ctx.last_timings
        print(json.dumps({"dest": name, "wall_ms": round(wall * 1e3, 2),
                          **{k: round(t[k] * 1e3, 3) for k in ("h2d", "pack", "get_lms", "lms_sort", "put_lms", "induce", "d2h")}}),
              flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
