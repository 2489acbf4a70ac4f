"""Seeded general-alphabet texts (bytes of any values) for tests/test_general.py and tests/golden/make_general_golden.py.

`UNIQUE`: cases where the reference's result is unique — no two suffixes share their first k symbols, or k >= n (the true
suffix array) — so the CUDA path, the oracle and the reference must agree bit for bit.  `TIED`: repetitive texts at a
small k, where the reference's own order among the tied suffixes is unspecified (an unstable std::sort,
kiss1_core.hpp:214-219): only the k-order property and the permutation are checked there."""
import numpy as np

import textgen as tg


def hashed(seed, n, alphabet):
    """symbol i = alphabet[splitmix64(seed' ^ i) mod sigma] (counter based: independent of numpy's Generator streams)"""
    alphabet = np.asarray(alphabet, np.uint8)
    if n == 0:
        return np.zeros(0, np.uint8)
    x = tg.sm64(tg.sm64(np.uint64(seed)) ^ np.arange(n, dtype=np.uint64))
    return alphabet[(x >> np.uint64(17)) % np.uint64(alphabet.size)]


AMINO = np.frombuffer(b"ACDEFGHIKLMNPQRSTVWY", np.uint8)      # the protein experiments' alphabet (experiment_a.sh:27-31)
ABCD = np.frombuffer(b"ABCD", np.uint8)                        # tests/kiss.cpp:53-87


def periodic(unit, n):
    u = np.frombuffer(unit, np.uint8)
    return np.tile(u, n // u.size + 1)[:n].copy()


def planted(seed, n, alphabet, unit_len, copies):
    """hashed background with `copies` exact copies of one unit"""
    S = hashed(seed, n, alphabet)
    unit = hashed(seed + 1, unit_len, alphabet)
    r = tg._Rng(seed)
    for _ in range(copies):
        p = r.below(n - unit_len - 8)
        S[p:p + unit_len] = unit
    return S


# "full order" for the cases below: k = 2^31 - 1.  (k = 2^32 - 1, the CLI's `-k -1`, is not usable with the reference's
# general kISS-1: `i + k` wraps in its small_cmp, kiss1_core.hpp:209-210, and the array comes out mis-ordered; kISS-2 and
# the port return the true suffix array there — tests/test_general.py::test_k_minus_one_is_the_true_suffix_array.)
FULL = 0x7FFFFFFF

# name -> (builder, k)
UNIQUE = {
    "empty": (lambda: np.zeros(0, np.uint8), 256),
    "one": (lambda: np.frombuffer(b"x", np.uint8), 256),
    "two_same": (lambda: np.frombuffer(b"zz", np.uint8), 256),
    "two_desc": (lambda: np.frombuffer(b"ba", np.uint8), 256),
    "banana": (lambda: np.frombuffer(b"banana", np.uint8), 256),
    "mississippi": (lambda: np.frombuffer(b"mississippi", np.uint8), 256),
    "abcd_1k": (lambda: hashed(1, 1000, ABCD), 256),
    "abcd_150k": (lambda: hashed(2, 150_000, ABCD), 256),
    "sigma1_full": (lambda: periodic(b"q", 777), FULL),
    "sigma2_full": (lambda: hashed(3, 5000, [7, 200]), FULL),
    "sigma3": (lambda: hashed(4, 30_000, [0, 1, 255]), 256),
    "sigma5": (lambda: hashed(5, 30_000, [10, 20, 30, 40, 50]), 256),
    "sigma16": (lambda: hashed(6, 50_000, np.arange(100, 116)), 256),
    "sigma17": (lambda: hashed(7, 50_000, np.arange(100, 117)), 256),
    "amino_100k": (lambda: hashed(8, 100_000, AMINO), 256),
    "amino_k8": (lambda: hashed(9, 20_000, AMINO), 8),
    "sigma64": (lambda: hashed(10, 60_000, np.arange(64)), 256),
    "sigma65": (lambda: hashed(11, 60_000, np.arange(191, 256)), 256),
    "bytes_all": (lambda: hashed(12, 200_000, np.arange(256)), 256),
    "text_ascii": (lambda: np.frombuffer((b"the quick brown fox jumps over the lazy dog. " * 40) + b"END", np.uint8), FULL),
    "periodic_full": (lambda: periodic(b"abcabcab", 4000), FULL),
    "planted_full": (lambda: planted(13, 60_000, AMINO, 3000, 12), FULL),
    "runs_full": (lambda: np.concatenate([hashed(14, 3000, AMINO), periodic(b"K", 5000), hashed(15, 3000, AMINO),
                                          periodic(b"K", 5000), hashed(16, 50, AMINO)]), FULL),
    "k_equals_n_minus_1": (lambda: periodic(b"xyxyz", 2000), 1999),
}

TIED = {
    "periodic_k16": (lambda: periodic(b"abcabcab", 4000), 16),
    "planted_k64": (lambda: planted(13, 60_000, AMINO, 3000, 12), 64),
    "runs_k256": (lambda: np.concatenate([hashed(14, 3000, AMINO), periodic(b"K", 5000), hashed(15, 3000, AMINO),
                                          periodic(b"K", 5000), hashed(16, 50, AMINO)]), 256),
    "sigma1_k3": (lambda: periodic(b"q", 777), 3),
    "bytes_planted_k32": (lambda: planted(17, 80_000, np.arange(256), 2000, 20), 32),
}


def k_order_violations(S, SA, k, chunk=1 << 20):
    """number of adjacent pairs with substr(SA[i-1], k) > substr(SA[i], k) — the predicate of tests/kiss.cpp:61-65
    (the shorter string first when one is a prefix of the other).  Vectorised: 8 symbols per step on the undecided pairs."""
    S = np.asarray(S, np.uint8)
    n = S.size
    SA = np.asarray(SA, np.int64)
    P = np.concatenate([S, np.zeros(8, np.uint8)])
    bad = 0
    for lo in range(1, SA.size, chunk):
        a, b = SA[lo - 1:min(SA.size, lo + chunk) - 1].copy(), SA[lo:lo + chunk].copy()
        und = np.arange(a.size)
        d = 0
        while und.size and d < k:
            ia, ib = a[und] + d, b[und] + d
            w = min(8, k - d)
            res = np.zeros(und.size, np.int8)      # -1: a < b decided, +1: a > b decided
            for t in range(w):
                live = res == 0
                ea, eb = ia + t >= n, ib + t >= n   # past the end: the shorter one is smaller
                ca, cb = P[np.minimum(ia + t, n)], P[np.minimum(ib + t, n)]
                lt = live & ((ea & ~eb) | (~ea & ~eb & (ca < cb)))
                gt = live & ((~ea & eb) | (~ea & ~eb & (ca > cb)))
                both = live & ea & eb
                res[lt | both] = -1
                res[gt] = 1
            bad += int((res == 1).sum())
            und = und[res == 0]
            d += w
    return bad
