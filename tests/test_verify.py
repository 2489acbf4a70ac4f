"""kiss_b200_verify_sa_dev: the predicate of the reference's own tests (tests/kiss.cpp:25-33,
REQUIRE(substr(sa[i-1], k) <= substr(sa[i], k))) plus "permutation of 0..n", evaluated on the device over the whole array.
The checker itself is checked here: zero on correct arrays (the CUDA path's and the oracle's), non-zero on every kind of
damage, and tolerant exactly where the predicate is (suffixes that agree on their first k bases may swap)."""
import numpy as np
import pytest

import oracle_lib as ol
import textgen as tg

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import kiss_b200
    return kiss_b200.Context(0)


def verify(ctx, S, SA, k):
    import torch
    dS = torch.zeros(S.size + 64, dtype=torch.uint8, device="cuda")
    dS[:S.size] = torch.from_numpy(np.array(S, dtype=np.uint8, copy=True))
    dSA = torch.from_numpy(np.ascontiguousarray(SA).view(np.int32)).cuda()
    torch.cuda.synchronize()
    return ctx.verify_sa_dev(dS.data_ptr(), S.size, dSA.data_ptr(), k)


CLEAN = dict(order_violations=0, duplicates=0, out_of_range=0, cut=0)


@pytest.mark.parametrize("name,gen,k", [
    ("uniform", lambda: tg.uniform(1, 100_000), 256),
    ("mixed_k32", lambda: tg.mixed(2, 80_000), 32),
    ("repeats_full", lambda: tg.planted_repeats(3, 60_000, unit_len=2000, copies=10), 0xFFFFFFFF),
    ("periodic_k2", lambda: tg.periodic(5000, [1, 0, 2]), 2),
    ("all_A", lambda: np.zeros(3000, np.uint8), 256),
    ("one", lambda: tg.uniform(4, 1), 256),
    ("empty", lambda: np.zeros(0, np.uint8), 256),
])
def test_correct_arrays_are_clean(ctx, name, gen, k):
    S = gen()
    assert verify(ctx, S, ol.oracle_kiss1(S, k), k) == CLEAN          # the oracle's array
    assert verify(ctx, S, ctx.suffix_array_dna(S, k), k) == CLEAN     # the CUDA path's


def test_every_kind_of_damage_is_counted(ctx):
    S = tg.uniform(5, 50_000)
    SA = ol.oracle_kiss1(S, 256)
    bad = SA.copy(); bad[[1000, 1001]] = bad[[1001, 1000]]            # two neighbours swapped
    assert verify(ctx, S, bad, 256)["order_violations"] >= 1
    bad = SA.copy(); bad[777] = bad[778]                               # one value twice (and one missing)
    v = verify(ctx, S, bad, 256)
    assert v["duplicates"] == 1
    bad = SA.copy(); bad[5] = S.size + 7                               # out of range
    assert verify(ctx, S, bad, 256)["out_of_range"] == 1
    bad = SA.copy(); bad[0], bad[1] = bad[1], bad[0]                   # the sentinel is not first
    v = verify(ctx, S, bad, 256)
    assert v["out_of_range"] >= 1 or v["order_violations"] >= 1
    assert verify(ctx, S, np.sort(SA), 256)["order_violations"] > 10_000   # text order is not suffix order


def test_tolerant_exactly_where_the_predicate_is(ctx):
    """suffixes that agree on their first k bases may stand in either order (<=), one base more and they may not"""
    S = np.concatenate([tg.uniform(6, 4000), tg.uniform(7, 300), tg.uniform(8, 4000), tg.uniform(7, 300), tg.uniform(9, 500)])
    k = 64
    SA = ol.oracle_kiss1(S, k)
    a, b = 4000, 8300                                                  # two copies of the same 300 bases
    ia, ib = int(np.nonzero(SA == a)[0][0]), int(np.nonzero(SA == b)[0][0])
    assert abs(ia - ib) == 1                                           # equal for 300 > k bases: neighbours
    sw = SA.copy(); sw[[ia, ib]] = sw[[ib, ia]]
    assert verify(ctx, S, sw, k) == CLEAN
    assert verify(ctx, S, sw, 0xFFFFFFFF)["order_violations"] == 1     # at full depth exactly one of the two orders is right


def test_shorter_suffix_first_at_the_text_end(ctx):
    S = np.zeros(40, np.uint8)                                         # all A: suffix i is a prefix of suffix j < i
    good = np.arange(40, -1, -1).astype(np.uint32)
    assert verify(ctx, S, good, 256) == CLEAN
    bad = good.copy(); bad[[10, 11]] = bad[[11, 10]]
    assert verify(ctx, S, bad, 256)["order_violations"] == 1
    assert verify(ctx, S, bad, 4) == CLEAN                              # (suffixes 30 and 29 agree on their first 4 bases)


def test_general_alphabet_arrays_verify_through_their_reduction(ctx):
    """a byte text over <= 4 symbols IS a DNA text after ranking: the general entry's result passes the DNA verifier"""
    import general_cases as gc
    S = gc.hashed(41, 60_000, gc.ABCD)
    SA = ctx.suffix_array(S, 256)
    assert verify(ctx, S - 65, SA, 256) == CLEAN


def verify_bytes(ctx, S, SA, k):
    import torch
    dS = torch.zeros(S.size + 16, dtype=torch.uint8, device="cuda")
    dS[:S.size] = torch.from_numpy(np.array(S, dtype=np.uint8, copy=True))
    dSA = torch.from_numpy(np.ascontiguousarray(SA).view(np.int32)).cuda()
    torch.cuda.synchronize()
    return ctx.verify_sa_dev(dS.data_ptr(), S.size, dSA.data_ptr(), k, byte_text=True)


def test_byte_text_verifier(ctx):
    """kiss_b200_verify_sa_bytes_dev: the same predicate over a general-alphabet text, agreeing with the host checker of
    tests/general_cases.py"""
    import general_cases as gc
    for name, (build, k) in list(gc.UNIQUE.items()) + list(gc.TIED.items()):
        S = build()
        SA = ctx.suffix_array(S, k)
        assert verify_bytes(ctx, S, SA, k) == CLEAN, name
    S = gc.hashed(51, 30_000, gc.AMINO)
    SA = ol.oracle_general(S, 256)
    assert verify_bytes(ctx, S, SA, 256) == CLEAN
    bad = SA.copy(); bad[[200, 201]] = bad[[201, 200]]
    v = verify_bytes(ctx, S, bad, 256)
    assert v["order_violations"] == gc.k_order_violations(S, bad, 256) >= 1
    bad = SA.copy(); bad[9] = bad[10]
    assert verify_bytes(ctx, S, bad, 256)["duplicates"] == 1
    # bytes above 127 compare unsigned
    S = np.array([200, 100, 200, 50, 255, 0], np.uint8)
    SA = ol.oracle_general(S, 256)
    assert verify_bytes(ctx, S, SA, 256) == CLEAN
