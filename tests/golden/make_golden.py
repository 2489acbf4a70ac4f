#!/usr/bin/env python
"""Generates tests/golden/kiss1_golden.json from the REFERENCE ITSELF (oracle/_ref/libkiss_ref.so = the unmodified
jhhung/kISS headers compiled by oracle/build_ref.sh).  Run in the build container, where /root/reference exists:

    python tests/golden/make_golden.py

Every vector = (generator call, k) -> n, m, FNV-1a of the reference's LMS list, sorted LMS list and suffix array,
plus the first/last entries, so that the GPU box (which has no /root/reference) can check the oracle, the .so that
travelled and the CUDA path against the reference's own outputs."""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as ol  # noqa: E402
import textgen as tg  # noqa: E402

CASES = [
    ("uniform", dict(seed=101, n=1000)),
    ("uniform", dict(seed=102, n=65_537)),
    ("uniform", dict(seed=103, n=1_000_000)),
    ("planted_repeats", dict(seed=104, n=500_000, unit_len=5000, copies=60)),
    ("planted_repeats", dict(seed=105, n=500_000, unit_len=3000, copies=40, mutate_every=250)),
    ("tandem", dict(seed=106, n=300_000)),
    ("mixed", dict(seed=107, n=250_000)),
    ("low_entropy", dict(seed=108, n=40_000, alphabet=2)),
]
KS = [256, 32, 125, 0xFFFFFFFF]


def gen(name, kw):
    S = getattr(tg, name)(**kw)
    if name == "low_entropy":
        S = np.concatenate([S, tg.uniform(kw["seed"] + 1, 1024)])
    return S


def main():
    assert ol.have_ref(), "build oracle/_ref first (make -C oracle)"
    out = []
    for name, kw in CASES:
        S = gen(name, kw)
        for k in KS:
            r = ol.ref_kiss1_stages(S, k=k, threads=3)
            out.append(dict(gen=name, args=kw, k=k, n=int(S.size), m=int(r["m"]),
                            count=[int(x) for x in r["count"]], lms_count=[int(x) for x in r["lms_count"]],
                            lms_fnv=f"{ol.fnv1a(r['lms']):016x}", sorted_fnv=f"{ol.fnv1a(r['sorted']):016x}",
                            sa_fnv=f"{ol.fnv1a(r['sa']):016x}", sa_head=[int(x) for x in r["sa"][:8]],
                            sa_tail=[int(x) for x in r["sa"][-8:]]))
    json.dump(dict(source="jhhung/kISS kiss1_suffix_array_dna via oracle/ref_harness.cpp", vectors=out),
              open(os.path.join(HERE, "kiss1_golden.json"), "w"), indent=1)
    print(len(out), "vectors")


if __name__ == "__main__":
    main()
