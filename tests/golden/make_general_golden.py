#!/usr/bin/env python
"""tests/golden/general_golden.json from the REFERENCE ITSELF (oracle/_ref: ref_kiss1_suffix_array / ref_kiss2_suffix_array,
the unmodified general-alphabet kiss1_suffix_array / kiss2_suffix_array): one FNV-1a digest of the suffix array per UNIQUE
case of tests/general_cases.py.  The script refuses a case whose result is not unique (two adjacent suffixes sharing their
first k symbols, k < n) or on which kISS-1 and kISS-2 disagree.

    python tests/golden/make_general_golden.py"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import general_cases as gc  # noqa: E402
import oracle_lib as ol  # noqa: E402


def shares_k(S, SA, k):
    """some adjacent pair of SA agrees on its first k symbols (both at least k long)"""
    n = S.size
    if k >= n:
        return False
    a, b = SA[:-1].astype(np.int64), SA[1:].astype(np.int64)
    ok = (a + k <= n) & (b + k <= n)
    a, b = a[ok], b[ok]
    alive = np.ones(a.size, bool)
    for t in range(k):
        alive[alive] &= S[a[alive] + t] == S[b[alive] + t]
        if not alive.any():
            return False
    return bool(alive.any())


out = {}
for name, (build, k) in gc.UNIQUE.items():
    S = build()
    threads = 1 if S.size < 100_000 else 4    # (the reference's multi-threaded general path crashes on small texts over large alphabets)
    r1 = ol.ref_general(S, k, threads, 0)
    r2 = ol.ref_general(S, k, 1, 1)
    if S.size == 1:   # kISS-1's general path returns {0, EMPTY} for a one-symbol text (its LMS list does not fit the
        r1 = r2       # n / 2 = 0 slots it reserved, kiss1_core.hpp:289-293); kISS-2 returns the suffix array {1, 0}
    assert (r1 == r2).all(), name
    assert not shares_k(S, r1, k), name
    assert (ol.oracle_general(S, k) == r1).all(), name
    out[name] = {"n": int(S.size), "k": int(k), "sigma": int(np.unique(S).size), "fnv1a": "%016x" % ol.fnv1a(r1)}
json.dump(out, open(os.path.join(HERE, "general_golden.json"), "w"), indent=1)
print(len(out), "vectors")
