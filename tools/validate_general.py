#!/usr/bin/env python
"""Full-size parity of the general-alphabet entry: kiss_b200_suffix_sort_u32 against the REFERENCE ITSELF
(oracle/_ref/libkiss_ref.so: the unmodified kiss1_suffix_array, all host threads), bit-exact over the whole suffix array, on
uniform texts (no two suffixes share k symbols, so the reference's result is unique).  `protein_100m` has n >= 10^8 and so
goes through the reference's 32-symbol block comparator (kiss1_core.hpp:170-206, 214-225).
Run on the GPU box:  python tools/validate_general.py [abcd_20m protein_100m bytes_50m]   (one JSON line each)"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import general_cases as gc  # noqa: E402
import kiss_b200  # noqa: E402
import oracle_lib as ol  # noqa: E402

CONFIGS = {
    "abcd_20m": dict(n=20_000_000, seed=41, alphabet=gc.ABCD, k=256),          # tests/kiss.cpp:70-87 "general large"
    "protein_100m": dict(n=100_000_000, seed=42, alphabet=gc.AMINO, k=256),    # experiment_a.sh:27-31
    "bytes_50m": dict(n=50_000_000, seed=43, alphabet=np.arange(256), k=256),
}


def hashed_big(seed, n, alphabet, step=1 << 25):
    out = np.empty(n, np.uint8)
    alphabet = np.asarray(alphabet, np.uint8)
    s = gc.tg.sm64(np.uint64(seed))
    for lo in range(0, n, step):
        hi = min(n, lo + step)
        x = gc.tg.sm64(s ^ np.arange(lo, hi, dtype=np.uint64))
        out[lo:hi] = alphabet[(x >> np.uint64(17)) % np.uint64(alphabet.size)]
    return out


def main():
    names = sys.argv[1:] or list(CONFIGS)
    ctx = kiss_b200.Context(0)
    threads = os.cpu_count() or 8
    for name in names:
        cfg = CONFIGS[name]
        S = hashed_big(cfg["seed"], cfg["n"], cfg["alphabet"])
        assert (S[:1000] == gc.hashed(cfg["seed"], 1000, cfg["alphabet"])).all()
        ctx.suffix_array(S[:100_000], cfg["k"])          # warm-up
        best = None
        for _ in range(2):
            t0 = time.perf_counter()
            SA = ctx.suffix_array(S, cfg["k"])
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        t = ctx.last_timings
        t0 = time.perf_counter()
        ref = ol.ref_general(S, cfg["k"], threads, 0)
        ref_s = time.perf_counter() - t0
        import torch
        dS = torch.from_numpy(S).cuda()
        dSA = torch.from_numpy(SA.view(np.int32)).cuda()
        torch.cuda.synchronize()
        allpairs = ctx.verify_sa_dev(dS.data_ptr(), cfg["n"], dSA.data_ptr(), cfg["k"], byte_text=True)
        del dS, dSA
        rec = dict(config=name, k_order_all_pairs=allpairs, n=cfg["n"], sigma=int(len(cfg["alphabet"])), k=cfg["k"], gpu_host_call_s=round(best, 4),
                   gpu_device_s=round(t["total_device"], 4), ref_s=round(ref_s, 3), ref_threads=threads,
                   bit_exact=bool((SA == ref).all()), mismatches=int((SA != ref).sum()),
                   sa0_is_n=bool(SA[0] == cfg["n"]), fnv1a="%016x" % ol.fnv1a(SA))
        print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
