"""Deterministic DNA test-text generators (numpy, version-independent).

The uniform stream is the same counter-based generator the product's synthetic-input kernels use
(kiss_b200/csrc/synth.cuh): base i = (splitmix64(seed' ^ (i >> 5)) >> 2*(i & 31)) & 3 with
seed' = splitmix64(seed).  Everything else here builds adversarial small/medium inputs for the
parity tests: planted exact repeats, tandem arrays, homopolymer runs, low-entropy alphabets.
"""
import sys

import numpy as np

_M = np.uint64(0xFFFFFFFFFFFFFFFF)


def sm64(z):
    z = np.asarray(z, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = z + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def uniform(seed, n, start=0):
    """codes 0..3, positions start .. start+n of the uniform stream `seed`."""
    if n == 0:
        return np.zeros(0, np.uint8)
    s = sm64(np.uint64(seed))
    i = np.arange(start, start + n, dtype=np.uint64)
    x = sm64(s ^ (i >> np.uint64(5)))
    return ((x >> (np.uint64(2) * (i & np.uint64(31)))) & np.uint64(3)).astype(np.uint8)


def uniform_big(seed, n, step=1 << 27):
    """uniform(seed, n) for large n: one hash per 32-base word instead of one per base, in chunks (same values)"""
    out = np.empty(n, np.uint8)
    s = sm64(np.uint64(seed))
    for lo in range(0, n, step):          # step is a multiple of 32
        hi = min(n, lo + step)
        nw = (hi - lo + 31) // 32
        x = sm64(s ^ np.arange(lo // 32, lo // 32 + nw, dtype=np.uint64))
        b = x.view(np.uint8) if sys.byteorder == "little" else x.byteswap().view(np.uint8)   # byte j: bases 4j .. 4j+3
        q = np.empty((b.size, 4), np.uint8)
        for t in range(4):
            q[:, t] = (b >> (2 * t)) & 3
        out[lo:hi] = q.reshape(-1)[:hi - lo]
    return out


class _Rng:
    """tiny splitmix-based scalar RNG so fixtures never depend on numpy's Generator streams"""

    def __init__(self, seed):
        self.s = int(sm64(np.uint64(seed ^ 0xA5A5A5A5)))
        self.c = 0

    def u64(self):
        self.c += 1
        return int(sm64(np.uint64((self.s + self.c * 0x632BE59BD9B4E019) & 0xFFFFFFFFFFFFFFFF)))

    def below(self, n):
        return self.u64() % n

    def between(self, a, b):
        return a + self.below(b - a + 1)


def planted_repeats(seed, n, unit_len=5000, copies=100, mutate_every=0, tail=512):
    """uniform background with `copies` copies of one unit pasted at random places; with
    mutate_every > 0 each copy gets one substitution every ~mutate_every bases.  The last `tail`
    bases stay uniform (keeps the reference's comparator a strict weak order)."""
    S = uniform(seed, n).copy()
    r = _Rng(seed)
    unit = uniform(seed + 1000003, unit_len)
    if n <= unit_len + tail + 1:
        return S
    for _ in range(copies):
        p = r.below(n - unit_len - tail)
        S[p:p + unit_len] = unit
        if mutate_every:
            q = p + r.below(mutate_every)
            while q < p + unit_len:
                S[q] = (S[q] + 1 + r.below(3)) % 4
                q += 1 + r.below(2 * mutate_every)
    return S


def tandem(seed, n, arrays=20, max_unit=50, max_len=20000, tail=512):
    """uniform background with tandem arrays (unit length 1..max_unit; unit length 1 = homopolymer run)"""
    S = uniform(seed, n).copy()
    r = _Rng(seed)
    for _ in range(arrays):
        ul = r.between(1, max_unit)
        L = r.between(ul, max(ul, min(max_len, max(1, (n - tail) // 4))))
        if n - tail - L <= 0:
            continue
        p = r.below(n - tail - L)
        unit = uniform(seed + 7 * (_ + 1), ul)
        reps = -(-L // ul)
        S[p:p + L] = np.tile(unit, reps)[:L]
    return S


def many_runs(seed, n, runs, min_len, max_len, tail=512):
    """uniform background with `runs` homopolymer runs (all four symbols) of min_len..max_len bases: the induction
    chains inside a bucket then shrink from ~runs entries per wave to a handful (the on-device tail of the wave
    kernels: tile path above 512 entries, one entry per thread up to 512, one warp up to 32)"""
    S = uniform(seed, n).copy()
    r = _Rng(seed)
    for _ in range(runs):
        L = r.between(min_len, max_len)
        p = r.below(n - tail - L)
        S[p:p + L] = r.below(4)
    return S


def low_entropy(seed, n, alphabet=2):
    """text over the first `alphabet` symbols only (long ties everywhere)"""
    return (uniform(seed, n) % alphabet).astype(np.uint8)


def periodic(n, unit):
    unit = np.asarray(unit, np.uint8)
    return np.tile(unit, -(-n // len(unit)))[:n].copy()


def mixed(seed, n):
    """repeats + tandem + runs in one text"""
    S = planted_repeats(seed, n, unit_len=max(8, min(3000, n // 8)), copies=12, mutate_every=400,
                        tail=min(512, n // 4))
    T = tandem(seed + 1, n, arrays=6, max_unit=12, max_len=max(2, n // 10), tail=min(512, n // 4))
    mask = uniform(seed + 2, n // 64 + 1).repeat(64)[:n] == 0
    S[mask] = T[mask]
    return S


def chm13_like(seed, n, base=None, families=64, tandem_arrays=2000, segdups=500, hp_runs=2000):
    """uniform background + interspersed repeat families (copies mutated 2-15 %, ~35 % of n),
    tandem arrays (unit 2..2000, length 1 kb..1 Mb, 0-2 % divergence, ~10 %), exact segmental duplications
    (10-100 kb), homopolymer runs <= 64; the last 4096 bases stay uniform.  `base` (uint8, modified in place)
    may hold the uniform stream already (e.g. generated on the GPU).  Sizes scale down with n."""
    S = uniform(seed, n).copy() if base is None else base
    if n < 100_000:
        return S
    rng = np.random.default_rng(seed)
    lim = n - 4096
    scale = min(1.0, n / 3.1e9)

    def mutate(seg, rate):
        k = rng.binomial(seg.size, rate)
        if k:
            p = rng.integers(0, seg.size, k)
            seg[p] = (seg[p] + rng.integers(1, 4, k).astype(np.uint8)) & 3

    # (a) interspersed families
    target = int(0.35 * n)
    lens = [300, 1000, 6000]
    cons = [rng.integers(0, 4, lens[f % 3]).astype(np.uint8) for f in range(families)]
    done = 0
    while done < target:
        f = int(rng.integers(0, families))
        c = cons[f].copy()
        mutate(c, float(rng.uniform(0.02, 0.15)))
        p = int(rng.integers(0, lim - c.size))
        S[p:p + c.size] = c
        done += c.size
    # (b) tandem arrays
    target = int(0.10 * n)
    done = 0
    na = max(4, int(tandem_arrays * scale))
    for _ in range(na):
        if done >= target:
            break
        ul = int(np.exp(rng.uniform(np.log(2), np.log(2000))))
        L = int(np.exp(rng.uniform(np.log(1000), np.log(1_000_000))))
        L = min(L, lim // 8)
        unit = rng.integers(0, 4, ul).astype(np.uint8)
        arr = np.tile(unit, L // ul + 1)[:L]
        mutate(arr, float(rng.uniform(0.0, 0.02)))
        p = int(rng.integers(0, lim - L))
        S[p:p + L] = arr
        done += L
    # (c) exact segmental duplications
    for _ in range(max(2, int(segdups * scale))):
        L = int(rng.integers(10_000, 100_000))
        a, b = int(rng.integers(0, lim - L)), int(rng.integers(0, lim - L))
        S[b:b + L] = S[a:a + L].copy()
    # (d) homopolymer runs capped at 64
    for _ in range(max(4, int(hp_runs * scale))):
        L = int(rng.integers(8, 65))
        p = int(rng.integers(0, lim - L))
        S[p:p + L] = int(rng.integers(0, 4))
    return S


def near_identical_copies(seed, unit_len, copies, rate, tail=600):
    """`copies` copies of one random unit, each base substituted with probability `rate`, then `tail` uniform bases:
    LMS suffixes at the same offset of every copy share long prefixes (tie groups of size `copies`: a warp team up
    to 256, a block team up to 2048, the global pivot rounds above) and differ only where a copy was mutated."""
    rng = np.random.default_rng(seed)
    unit = rng.integers(0, 4, unit_len).astype(np.uint8)
    S = np.tile(unit, copies)
    k = rng.binomial(S.size, rate)
    if k:
        p = rng.integers(0, S.size, k)
        S[p] = (S[p] + rng.integers(1, 4, k).astype(np.uint8)) & 3
    return np.concatenate([S, uniform(seed + 1, tail)])
