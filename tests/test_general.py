"""General-alphabet path: kiss_b200_suffix_sort_u32 against the reference's kiss1_suffix_array /
kiss2_suffix_array (kiss1_core.hpp:270-311, kiss2_core.hpp:898-906).

CPU tests pin oracle/general_oracle.c to the reference (live through oracle/_ref when present, and through
tests/golden/general_golden.json, generated from the reference by tests/golden/make_general_golden.py).  GPU tests run the
CUDA path through the C ABI: bit-exact on every case whose reference result is unique, the k-order property of the
reference's own tests (tests/kiss.cpp:53-87) where it is not."""
import json
import os

import numpy as np
import pytest

import general_cases as gc
import oracle_lib as ol

GOLD = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "general_golden.json")))
UNIQUE = sorted(gc.UNIQUE)
TIED = sorted(gc.TIED)


def test_every_unique_case_has_a_golden_vector():
    assert sorted(GOLD) == UNIQUE


# ------------------------------------------------------------------------------------------ the oracle
@pytest.mark.parametrize("name", UNIQUE)
def test_oracle_matches_the_reference_golden(name):
    build, k = gc.UNIQUE[name]
    S = build()
    assert S.size == GOLD[name]["n"] and k == GOLD[name]["k"]
    assert "%016x" % ol.fnv1a(ol.oracle_general(S, k)) == GOLD[name]["fnv1a"]


@pytest.mark.skipif(not ol.have_ref(), reason="oracle/_ref not built")
@pytest.mark.parametrize("name", [n for n in UNIQUE if gc.UNIQUE[n][0]().size != 1])
def test_oracle_matches_the_live_reference(name):
    build, k = gc.UNIQUE[name]
    S = build()
    want = ol.oracle_general(S, k)
    assert (ol.ref_general(S, k, 1, 0) == want).all()
    assert (ol.ref_general(S, k, 1, 1) == want).all()


@pytest.mark.parametrize("name", ["banana", "mississippi", "periodic_full", "text_ascii", "sigma1_full"])
def test_oracle_full_order_is_the_naive_suffix_array(name):
    S = gc.UNIQUE[name][0]()
    n = S.size
    want = np.array([n] + sorted(range(n), key=lambda i: bytes(S[i:])), np.uint32)
    assert (ol.oracle_general(S, gc.FULL) == want).all()
    assert (ol.oracle_general(S, 0xFFFFFFFF) == want).all()


@pytest.mark.parametrize("name", TIED)
def test_oracle_is_k_ordered_on_tied_texts(name):
    build, k = gc.TIED[name]
    S = build()
    SA = ol.oracle_general(S, k)
    assert (np.sort(SA) == np.arange(S.size + 1)).all()
    assert gc.k_order_violations(S, SA, k) == 0


@pytest.mark.skipif(not ol.have_ref(), reason="oracle/_ref not built")
@pytest.mark.parametrize("name", TIED)
def test_reference_is_k_ordered_on_tied_texts(name):
    """what the reference's own tests assert, on texts where its order is not unique — the bar the CUDA path is held to"""
    build, k = gc.TIED[name]
    S = build()
    SA = ol.ref_general(S, k, 1, 0)
    assert (np.sort(SA) == np.arange(S.size + 1)).all()
    assert gc.k_order_violations(S, SA, k) == 0


def test_k_order_checker_finds_a_swap():
    S = gc.hashed(21, 5000, gc.AMINO)
    SA = ol.oracle_general(S, 256)
    assert gc.k_order_violations(S, SA, 256) == 0
    SA[[100, 101]] = SA[[101, 100]]
    assert gc.k_order_violations(S, SA, 256) >= 1


def test_block_comparator_restatement():
    """oracle_general_less_big (kiss1_core.hpp:170-206, used by the reference from 10^8 symbols on) agrees with the plain
    prefix order wherever two suffixes differ inside the compared depth"""
    S = gc.planted(22, 20_000, gc.ABCD, 500, 6)
    n, k = S.size, 64
    D = 32 * (k // 32 + 1)
    rng = np.random.default_rng(3)
    L = ol.oracle()
    for _ in range(3000):
        i, j = (int(x) for x in rng.integers(0, n, 2))
        if i == j:
            continue
        a, b = bytes(S[i:i + D]), bytes(S[j:j + D])
        if a != b and len(a) == D and len(b) == D:
            assert bool(L.oracle_general_less_big(S.ctypes.data, n, k, i, j)) == (a < b)
        elif a == b:
            assert bool(L.oracle_general_less_big(S.ctypes.data, n, k, i, j)) == (i < j)


# ------------------------------------------------------------------------------------------ the CUDA path
@pytest.fixture(scope="module")
def ctx():
    import kiss_b200
    return kiss_b200.Context(0)


@pytest.mark.gpu
@pytest.mark.parametrize("algo", [0, 1])
@pytest.mark.parametrize("name", UNIQUE)
def test_cuda_matches_oracle_and_golden(ctx, name, algo):
    build, k = gc.UNIQUE[name]
    S = build()
    SA = ctx.suffix_array(S, k, algo)
    assert SA.size == S.size + 1
    assert (SA == ol.oracle_general(S, k)).all()
    assert "%016x" % ol.fnv1a(SA) == GOLD[name]["fnv1a"]


@pytest.mark.gpu
@pytest.mark.parametrize("algo", [0, 1])
@pytest.mark.parametrize("name", TIED)
def test_cuda_is_k_ordered_on_tied_texts(ctx, name, algo):
    build, k = gc.TIED[name]
    S = build()
    SA = ctx.suffix_array(S, k, algo)
    assert (np.sort(SA) == np.arange(S.size + 1)).all()
    assert gc.k_order_violations(S, SA, k) == 0


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["text_ascii", "periodic_full", "runs_full", "planted_full"])
def test_k_minus_one_is_the_true_suffix_array(ctx, name):
    """k = 2^32 - 1 (`-k -1`): the true suffix array, what kISS-2 returns (the reference's general kISS-1 wraps there)"""
    S = gc.UNIQUE[name][0]()
    want = ol.oracle_general(S, 0xFFFFFFFF)
    for algo in (0, 1):
        assert (ctx.suffix_array(S, -1, algo) == want).all()


@pytest.mark.gpu
@pytest.mark.parametrize("lo,hi", [(100_000, 200_000), (10_000_000, 20_000_000)])
def test_the_references_own_general_tests(ctx, lo, hi):
    """tests/kiss.cpp:53-87 ("kISS-1 general", "kISS-1 general large"): symbols 'A'..'D', k = 256,
    REQUIRE(substr(sa[i-1], k) <= substr(sa[i], k)); the length is drawn once, seeded"""
    import kiss_b200
    n = lo + int(gc.tg.sm64(np.uint64(hi)) % np.uint64(hi - lo))
    S = gc.hashed(31, n, gc.ABCD)
    SA = kiss_b200.KISS1Sorter.get_suffix_array(S, 256)
    assert SA[0] == n and SA.size == n + 1
    assert gc.k_order_violations(S, SA, 256) == 0
    # uniform text: no two suffixes share 256 symbols, so this is also the true suffix array — one digest against the
    # DNA entry on the same codes (b = 1: the reduced text IS a DNA text)
    assert (SA == ctx.suffix_array_dna(S - 65, 256)).all()


@pytest.mark.gpu
def test_kiss2_sorter_general(ctx):
    import kiss_b200
    S = gc.hashed(32, 300_000, gc.AMINO)
    SA = kiss_b200.KISS2Sorter.get_suffix_array(S, 256)
    assert (SA == ol.oracle_general(S, 256)).all()


@pytest.mark.gpu
def test_device_resident_entry_and_digits(ctx):
    import torch
    S = gc.hashed(33, 400_000, gc.AMINO)
    d = torch.from_numpy(S).cuda()
    out = torch.empty(S.size + 1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    assert ctx.general_digits(d.data_ptr(), S.size) == (20, 3)
    ctx.suffix_array_dev(d.data_ptr(), S.size, out.data_ptr(), 256)
    got = out.cpu().numpy().view(np.uint32)
    assert (got == ol.oracle_general(S, 256)).all()
    for alphabet, want in [([5], (1, 1)), ([1, 2, 3, 4], (4, 1)), (np.arange(5), (5, 2)), (np.arange(16), (16, 2)),
                           (np.arange(17), (17, 3)), (np.arange(64), (64, 3)), (np.arange(65), (65, 4)),
                           (np.arange(256), (256, 4))]:
        t = torch.from_numpy(gc.hashed(34, 50_000, alphabet)).cuda()
        torch.cuda.synchronize()
        assert ctx.general_digits(t.data_ptr(), t.numel()) == want


@pytest.mark.gpu
def test_reduced_text_beyond_32_bit_indices_is_refused(ctx):
    """b * n must fit 32-bit suffix indices: 1.2 * 10^9 symbols of 256 kinds need 4.8 * 10^9 bases"""
    import kiss_b200
    import torch
    n = 1_200_000_000
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    d.random_(0, 256)
    out = torch.empty(16, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    with pytest.raises(kiss_b200.KissB200Error) as e:
        ctx.suffix_array_dev(d.data_ptr(), n, out.data_ptr(), 256)
    assert "32-bit" in str(e.value)


# ------------------------------------------------------------------------------------------ CPU fuzz of the oracle
@pytest.mark.parametrize("seed", range(6))
def test_oracle_fuzz_against_the_naive_suffix_array_and_the_reference(seed):
    """random small byte texts (1..6 distinct symbols: ties everywhere), full order: the restatement, the naive suffix
    array and — when built — the reference's kISS-2 agree (kISS-2: its general kISS-1 breaks on n = 1 and k = 2^32 - 1)"""
    r = gc.tg._Rng(7000 + seed)
    for _ in range(60):
        n = [1, 2, 3, 4, 7, 16, 33, 100, 257, 600][r.below(10)]
        sigma = 1 + r.below(6)
        vals = sorted({r.below(256) for _ in range(sigma)})
        S = np.array([vals[r.below(len(vals))] for _ in range(n)], np.uint8)
        want = np.array([n] + sorted(range(n), key=lambda i: bytes(S[i:])), np.uint32)
        assert (ol.oracle_general(S, gc.FULL) == want).all(), S.tolist()
        if ol.have_ref():
            assert (ol.ref_general(S, gc.FULL, 1, 1) == want).all(), S.tolist()
            # kISS-1's general driver reserves n / 2 slots for the LMS list (kiss1_core.hpp:289-293), one short when n is odd
            # and every other position is LMS ([1, 0, 2]: one LMS position + the sentinel in one slot): it then returns
            # EMPTY entries.  Wherever its array is a permutation at all it must be the suffix array.
            r1 = ol.ref_general(S, gc.FULL, 1, 0)
            if (np.sort(r1) == np.arange(n + 1)).all():
                assert (r1 == want).all(), S.tolist()
