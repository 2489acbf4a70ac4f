#!/usr/bin/env python
"""Golden checksums of the REFERENCE's suffix arrays at the BASELINE.json sizes (C1 48.8 Mbp uniform, C2 1 Gbp uniform,
C3 3.1 Gbp chm13-like), made with oracle/_ref/libkiss_ref.so (the unmodified kiss1_suffix_array_dna, all
host threads) in the build container:

    python tests/golden/make_full_size_golden.py [c1 c2 c3 ...]      -> tests/golden/full_size_golden.json

Two digests per array: FNV-1a over the u32 entries (what tests/oracle_lib.fnv1a computes) and the position-mixed sum
    sum_i splitmix64((i << 32) | SA[i])  mod 2^64
which is order-sensitive yet a plain sum, so the device computes it in one pass (kiss_b200_sa_checksum_dev) and bench.py
compares it with the value stored here after every run, outside the timed region."""
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as ol  # noqa: E402
import textgen as tg  # noqa: E402

CONFIGS = {
    "c1": dict(n=48_800_648, seed=1, kind="uniform", k=256),
    "c2": dict(n=1_000_000_000, seed=2, kind="uniform", k=256),
    "c3": dict(n=3_117_292_070, seed=3, kind="chm13", k=256),
    "c3_200m": dict(n=200_000_000, seed=3, kind="chm13", k=256),      # the size of the -m gpu test
}
OUT = os.path.join(HERE, "full_size_golden.json")


def text(cfg):
    S = tg.uniform_big(cfg["seed"], cfg["n"])
    if cfg["kind"] == "chm13":
        S = tg.chm13_like(cfg["seed"], cfg["n"], base=S)
    return S


def mixed_sum(sa, step=1 << 24):
    tot = np.uint64(0)
    with np.errstate(over="ignore"):
        for lo in range(0, sa.size, step):
            hi = min(sa.size, lo + step)
            x = (np.arange(lo, hi, dtype=np.uint64) << np.uint64(32)) | sa[lo:hi].astype(np.uint64)
            tot = tot + tg.sm64(x).sum(dtype=np.uint64)
    return int(tot)


def main(names):
    res = json.load(open(OUT)) if os.path.exists(OUT) else {}
    threads = os.cpu_count() or 1
    for name in names:
        cfg = CONFIGS[name]
        t0 = time.time()
        S = text(cfg)
        t1 = time.time()
        SA = np.empty(cfg["n"] + 1, np.uint32)
        sec = ol.ref().ref_kiss1_suffix_array_dna(S.ctypes.data, S.size, cfg["k"], threads, SA.ctypes.data)
        assert int(SA[0]) == cfg["n"]
        res[name] = dict(cfg, text_sha256_head=ol.sha_bytes(S[:1 << 20]), sa_fnv=f"{ol.fnv1a(SA):016x}",
                         sa_mixed_sum=f"{mixed_sum(SA):016x}", reference_seconds=round(sec, 2), reference_threads=threads)
        print(name, res[name], f"text {t1 - t0:.0f} s", flush=True)
        del S, SA
        with open(OUT, "w") as f:
            json.dump(res, f, indent=1)


if __name__ == "__main__":
    main(sys.argv[1:] or ["c1", "c3_200m", "c2", "c3"])
