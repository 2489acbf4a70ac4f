#!/usr/bin/env python
"""Stage timings of repeated host-API calls on one of tools/validate_large.py's configurations (debug aid)."""
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import kiss_b200  # noqa: E402
from validate_large import CONFIGS, text  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c3small"
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    ctx = kiss_b200.Context(0)
    cfg = CONFIGS[name]
    S = text(ctx, cfg)
    for r in range(reps):
        ctx.suffix_array_dna(S, cfg["k"])
        t = ctx.last_timings
        print(json.dumps({"rep": r, **{k: round(t[k] * 1e3, 3) for k in
                                      ("h2d", "pack", "get_lms", "lms_sort", "put_lms", "induce", "d2h", "total_device")},
                          "launches": t["kernel_launches"]}), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
