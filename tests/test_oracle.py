"""CPU tests: the plain-C restatement (oracle/kiss_oracle.c) against the reference itself
(oracle/_ref/libkiss_ref.so, when present) stage by stage, and against naive suffix sorting."""
import numpy as np
import pytest

import oracle_lib as ol
import textgen as tg

needs_ref = pytest.mark.skipif(not ol.have_ref(), reason="oracle/_ref/libkiss_ref.so not built (no /root/reference)")

CASES = [
    ("uniform", lambda: tg.uniform(1, 1)),
    ("uniform", lambda: tg.uniform(2, 2)),
    ("uniform", lambda: tg.uniform(3, 7)),
    ("uniform", lambda: tg.uniform(4, 100)),
    ("uniform", lambda: tg.uniform(5, 1000)),
    ("uniform", lambda: tg.uniform(6, 4097)),
    ("uniform", lambda: tg.uniform(7, 200_000)),
    ("repeats", lambda: tg.planted_repeats(8, 300_000, unit_len=3000, copies=40)),
    ("repeats_mut", lambda: tg.planted_repeats(9, 300_000, unit_len=3000, copies=40, mutate_every=300)),
    ("tandem", lambda: tg.tandem(10, 200_000)),
    ("low_entropy", lambda: np.concatenate([tg.low_entropy(11, 50_000), tg.uniform(12, 1024)])),
    ("mixed", lambda: tg.mixed(13, 150_000)),
    ("all_A", lambda: np.zeros(500, np.uint8)),
    ("all_T", lambda: np.full(500, 3, np.uint8)),
    ("periodic_CA", lambda: np.concatenate([tg.periodic(5000, [1, 0]), tg.uniform(14, 700)])),
]


@needs_ref
@pytest.mark.parametrize("k", [256, 32, 125, 124, 2, 0xFFFFFFFF])
@pytest.mark.parametrize("name,gen", CASES, ids=[f"{c[0]}{i}" for i, c in enumerate(CASES)])
def test_restatement_matches_reference_stage_by_stage(name, gen, k):
    S = gen()
    # period-2 text makes a reference thread's chunk exactly half LMS; its branch-free
    # `SA[ptr] = i` (kiss_common.hpp:524) then writes one slot past its own output area into the
    # next thread's (a data race in the reference, seen with threads=3) — run that case on 1 thread.
    r = ol.ref_kiss1_stages(S, k=k, threads=1 if name == "periodic_CA" else 3)
    o = ol.oracle_kiss1(S, k=k, stages=True)
    assert o["m"] == r["m"]
    assert np.array_equal(o["count"], r["count"])
    assert np.array_equal(o["lms_count"], r["lms_count"])
    assert np.array_equal(o["lms"], r["lms"])
    assert np.array_equal(o["sorted"], r["sorted"])
    assert np.array_equal(o["placed"], r["placed"])
    assert np.array_equal(o["sa"], r["sa"])


@needs_ref
def test_reference_is_thread_count_invariant():
    # precondition for bit-exact parity
    S = tg.planted_repeats(21, 400_000, unit_len=4000, copies=60)
    a = ol.ref_kiss1(S, 256, threads=1)
    for t in (3, 8):
        assert np.array_equal(a, ol.ref_kiss1(S, 256, threads=t))


@pytest.mark.parametrize("n", [0, 1, 2, 3, 10, 64, 333])
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_unbounded_k_is_the_true_suffix_array(n, seed):
    S = tg.uniform(seed, n) if seed != 3 else tg.low_entropy(seed, n, 2)
    sa = ol.oracle_kiss1(S, k=0xFFFFFFFF)
    assert np.array_equal(sa, ol.naive_suffix_array(S))


def test_reference_test_property_k_order():
    # the only assertion the reference's own tests make (tests/kiss.cpp:26-28): adjacent suffixes
    # are non-decreasing in their first k bases; plus what it forgets: a permutation, SA[0] = n.
    S = tg.planted_repeats(31, 120_000, unit_len=2000, copies=30)
    k = 64
    sa = ol.oracle_kiss1(S, k=k)
    n = S.size
    assert sa[0] == n
    assert np.array_equal(np.sort(sa), np.arange(n + 1, dtype=np.uint32))
    b = bytes(S.tolist())
    prev = b[sa[0]:sa[0] + k]
    for i in range(1, n + 1):
        cur = b[sa[i]:sa[i] + k]
        assert prev <= cur
        prev = cur


def test_get_lms_definition():
    S = tg.mixed(41, 5000)
    lms, cnt, lcnt = ol.oracle_get_lms(S)
    n = S.size
    # brute-force types
    t = np.zeros(n + 1, np.int8)  # 1 = S
    t[n] = 1
    for i in range(n - 2, -1, -1):
        t[i] = 1 if (S[i] < S[i + 1] or (S[i] == S[i + 1] and t[i + 1] == 1)) else 0
    exp = [i for i in range(1, n) if t[i] == 1 and t[i - 1] == 0] + [n]
    assert lms.tolist() == exp
    assert cnt.tolist() == np.bincount(S, minlength=4).tolist()
    assert lcnt.tolist() == np.bincount(S[np.array(exp[:-1], dtype=np.int64)], minlength=4).tolist()
