"""numpy model of the GPU algorithm (NOT the oracle, NOT the product): the same decomposition the
CUDA kernels use — valley-rule LMS detection, word-wise D(k) keys + position for the long LMS
suffixes with comparator insertion of the few short ones, cursor-based wave induction with
pure character predicates — so that the *design* is checked against the oracle on the CPU before
the kernels are.  tests/test_model.py runs it.
"""
import numpy as np

import oracle_lib as ol

STRIDE = 125


def depth_D(k):
    """D(k) = 125*(k//125 + 1): bases the reference's block loop compares (kiss1_core.hpp:94-120)"""
    return STRIDE * (k // STRIDE + 1)


def lms_positions(S):
    """valley rule: i is LMS iff S[i-1] > S[i] and the run of S[i] starting at i is followed by a
    larger symbol (a run reaching the end of the text is L-type)."""
    n = S.size
    if n < 2:
        return np.zeros(0, np.int64)
    # s[i] = 1 iff type(i) = S ; right-to-left carry through equal runs
    lt = np.zeros(n, bool)
    eq = np.zeros(n, bool)
    lt[:-1] = S[:-1] < S[1:]
    eq[:-1] = S[:-1] == S[1:]
    s = np.zeros(n, bool)
    carry = False
    for i in range(n - 1, -1, -1):
        carry = lt[i] or (eq[i] and carry)
        s[i] = carry
    gt_prev = np.zeros(n, bool)
    gt_prev[1:] = S[:-1] > S[1:]
    return np.nonzero(s & gt_prev)[0]


def sort_lms(S, lms, k):
    """order of the LMS suffixes (sentinel excluded): long ones (>= D bases left) by their first D
    bases then position via a stable word-wise refinement; short ones by comparator insertion."""
    n = S.size
    D = depth_D(k)
    unbounded = D >= (1 << 32)
    lms = np.asarray(lms, np.int64)
    long_mask = (n - lms) >= D if not unbounded else np.zeros(lms.size, bool)
    longs = lms[long_mask]
    shorts = lms[~long_mask]
    # long: keys of D bases, compared word by word (32 bases); python bytes give the same order
    b = bytes(S.tolist())
    order = sorted(longs.tolist(), key=lambda p: (b[p:p + D], p))
    # shorts: insert with the reference comparator (binary search; consistent whenever the
    # comparator is a strict weak order)
    import functools
    lib = ol.oracle()
    Sp = np.ascontiguousarray(S)

    def pre10(i):  # the reference buckets by the zero-padded first 10 bases before it ever
        return b[i:i + 10].ljust(10, b"\0")  # calls its comparator (kiss1_core.hpp:41-83)

    def less(i, j):
        pi, pj = pre10(i), pre10(j)
        if pi != pj:
            return pi < pj
        return bool(lib.oracle_lms_less(Sp.ctypes.data, n, k & 0xFFFFFFFF, int(i), int(j)))

    shorts_sorted = sorted(shorts.tolist(), key=functools.cmp_to_key(lambda a, c: -1 if less(a, c) else (1 if less(c, a) else 0)))
    ins = []
    for sft in shorts_sorted:
        lo, hi = 0, len(order)
        while lo < hi:  # first index whose element is not less than sft
            mid = (lo + hi) // 2
            if less(order[mid], sft):
                lo = mid + 1
            else:
                hi = mid
        ins.append(lo)
    out = []
    si = 0
    for idx, p in enumerate(order):
        while si < len(ins) and ins[si] <= idx:
            out.append(shorts_sorted[si])
            si += 1
        out.append(p)
    out.extend(shorts_sorted[si:])
    return np.array(out, np.int64)


def induce_waves(S, sorted_lms):
    """cursor-based wave induction.  sorted_lms excludes the sentinel.  Returns (SA, n_waves)."""
    n = S.size
    S = S.astype(np.int64)
    cnt = np.bincount(S, minlength=4)
    B = np.concatenate([[1], 1 + np.cumsum(cnt)])  # bucket starts B[c], ends B[c+1]
    SA = np.full(n + 1, -1, np.int64)
    SA[0] = n
    lms_c = S[sorted_lms] if len(sorted_lms) else np.zeros(0, np.int64)
    lms_count = np.bincount(lms_c, minlength=4)
    off = 0
    for c in range(4):
        SA[B[c + 1] - lms_count[c]:B[c + 1]] = sorted_lms[off:off + lms_count[c]]
        off += lms_count[c]
    waves = 0

    def run_wave(beg, end, pred, cursors, down):
        nonlocal waves
        waves += 1
        idx = np.arange(beg, end) if not down else np.arange(end - 1, beg - 1, -1)
        v = SA[idx]
        assert (v >= 0).all()
        v = v[v > 0]
        u = v - 1
        cu = S[u]
        keep = pred(cu)
        u, cu = u[keep], cu[keep]
        for d in range(4):
            w = u[cu == d]
            if w.size == 0:
                continue
            if not down:
                SA[cursors[d]:cursors[d] + w.size] = w
                cursors[d] += w.size
            else:
                SA[cursors[d] - w.size:cursors[d]] = w[::-1]
                cursors[d] -= w.size

    # ---- L pass: slot 0, then per bucket: L-region waves, LMS segment
    Lc = B[:4].copy()
    run_wave(0, 1, lambda cu: np.ones(cu.shape, bool), Lc, False)
    for c in range(4):
        p = B[c]
        while p < Lc[c]:
            q = Lc[c]
            run_wave(p, q, lambda cu, c=c: cu >= c, Lc, False)
            p = q
        if lms_count[c]:
            run_wave(B[c + 1] - lms_count[c], B[c + 1], lambda cu: np.ones(cu.shape, bool), Lc, False)
    nL = Lc - B[:4]
    # ---- S pass: per bucket from the top: S-region waves (downwards), then the L region
    Sc = B[1:5].copy()
    for c in range(3, -1, -1):
        p = B[c + 1]
        while Sc[c] < p:
            q = Sc[c]
            run_wave(q, p, lambda cu, c=c: cu <= c, Sc, True)
            p = q
        if nL[c] and c > 0:
            run_wave(B[c], B[c] + nL[c], lambda cu, c=c: cu < c, Sc, True)
    assert (Sc == B[:4] + nL).all()
    return SA.astype(np.uint32), waves


def model_suffix_array(S, k=256):
    S = np.ascontiguousarray(S, np.uint8)
    n = S.size
    if n == 0:
        return np.zeros(1, np.uint32)
    lms = lms_positions(S)
    srt = sort_lms(S, lms, k)
    return induce_waves(S, srt)[0]
