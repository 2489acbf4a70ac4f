"""Deterministic DNA test-text generators (numpy, version-independent).

The uniform stream is the same counter-based generator the product's synthetic-input kernels use
(kiss_b200/csrc/synth.cuh): base i = (splitmix64(seed' ^ (i >> 5)) >> 2*(i & 31)) & 3 with
seed' = splitmix64(seed).  Everything else here builds adversarial small/medium inputs for the
parity tests: planted exact repeats, tandem arrays, homopolymer runs, low-entropy alphabets.
"""
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
