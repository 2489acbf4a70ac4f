"""GPU parity tests (run on the B200 box: `pytest -m gpu`): the CUDA path through the C ABI against the
oracle (oracle/kiss_oracle.c, pinned to the reference in tests/test_oracle.py) stage by stage."""
import numpy as np
import pytest

import oracle_lib as ol
import textgen as tg

pytestmark = pytest.mark.gpu

KS = [256, 32, 125, 124, 2, 0xFFFFFFFF]

CASES = [
    ("uniform1", lambda: tg.uniform(1, 1)),
    ("uniform2", lambda: tg.uniform(2, 2)),
    ("uniform7", lambda: tg.uniform(3, 7)),
    ("uniform100", lambda: tg.uniform(4, 100)),
    ("uniform1000", lambda: tg.uniform(5, 1000)),
    ("uniform4097", lambda: tg.uniform(6, 4097)),
    ("uniform200k", lambda: tg.uniform(7, 200_000)),
    ("repeats", lambda: tg.planted_repeats(8, 300_000, unit_len=3000, copies=40)),
    ("repeats_mut", lambda: tg.planted_repeats(9, 300_000, unit_len=3000, copies=40, mutate_every=300)),
    ("tandem", lambda: tg.tandem(10, 200_000)),
    ("low_entropy", lambda: np.concatenate([tg.low_entropy(11, 50_000), tg.uniform(12, 1024)])),
    ("mixed", lambda: tg.mixed(13, 150_000)),
    ("all_A", lambda: np.zeros(500, np.uint8)),
    ("all_T", lambda: np.full(500, 3, np.uint8)),
    ("periodic_CA", lambda: np.concatenate([tg.periodic(5000, [1, 0]), tg.uniform(14, 700)])),
    ("periodic_big", lambda: np.concatenate([tg.periodic(40_000, [2, 0, 1]), tg.uniform(15, 900)])),
    ("homopolymers", lambda: tg.tandem(16, 120_000, arrays=12, max_unit=1, max_len=9000)),
    # tie groups of near-identical copies: 40 (warp team), 200 (warp team, 8 slots per lane), 1500 (block team),
    # 5000 (pivot rounds in global memory); sparse and dense mutations, exact copies
    ("copies40", lambda: tg.near_identical_copies(17, 700, 40, 0.01)),
    ("copies200", lambda: tg.near_identical_copies(18, 500, 200, 0.004)),
    ("copies1500", lambda: tg.near_identical_copies(19, 150, 1500, 0.005)),
    ("copies5000", lambda: tg.near_identical_copies(20, 60, 5000, 0.003)),
    ("copies5000_dense", lambda: tg.near_identical_copies(21, 90, 5000, 0.08)),
    ("copies3000_exact", lambda: tg.near_identical_copies(22, 70, 3000, 0.0)),
    # hundreds / thousands of parallel homopolymer runs: waves of 33..512 entries (one entry per thread in the
    # on-device tail) and the hand-over from the tile path above 512 entries down to the single-warp path
    ("runs300", lambda: tg.many_runs(23, 200_000, 300, 20, 300)),
    ("runs3000", lambda: tg.many_runs(24, 600_000, 3000, 5, 120)),
]
IDS = [c[0] for c in CASES]


@pytest.fixture(scope="module")
def ctx():
    import kiss_b200
    c = kiss_b200.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("name,gen", CASES, ids=IDS)
def test_get_lms(ctx, name, gen):
    S = gen()
    lms, cnt, lcnt = ctx.get_lms(S)
    o_lms, o_cnt, o_lcnt = ol.oracle_get_lms(S)
    assert np.array_equal(lms, o_lms)
    assert cnt.tolist() == o_cnt.tolist()
    assert lcnt.tolist() == o_lcnt.tolist()


@pytest.mark.parametrize("k", KS)
@pytest.mark.parametrize("name,gen", CASES, ids=IDS)
def test_sorted_lms(ctx, name, gen, k):
    S = gen()
    o = ol.oracle_kiss1(S, k=k, stages=True)
    got = ctx.sorted_lms(S, k)
    assert got.size == o["m"]
    bad = np.nonzero(got != o["sorted"])[0]
    assert bad.size == 0, f"{bad.size} mismatches, first at {bad[:5]}: got {got[bad[:5]]} want {o['sorted'][bad[:5]]}"


@pytest.mark.parametrize("name,gen", CASES, ids=IDS)
def test_put_lms(ctx, name, gen):
    S = gen()
    o = ol.oracle_kiss1(S, k=256, stages=True)
    assert np.array_equal(ctx.put_lms(S, 256), o["placed"])


@pytest.mark.parametrize("name,gen", CASES, ids=IDS)
def test_induce_from_oracle_sorted_lms(ctx, name, gen):
    S = gen()
    o = ol.oracle_kiss1(S, k=256, stages=True)
    got = ctx.induce_from_sorted_lms(S, o["sorted"])
    bad = np.nonzero(got != o["sa"])[0]
    assert bad.size == 0, f"{bad.size} mismatches, first at {bad[:5]}"


@pytest.mark.parametrize("k", KS)
@pytest.mark.parametrize("name,gen", CASES, ids=IDS)
def test_full_suffix_array(ctx, name, gen, k):
    S = gen()
    want = ol.oracle_kiss1(S, k=k)
    got = ctx.suffix_array_dna(S, k)
    bad = np.nonzero(got != want)[0]
    assert bad.size == 0, f"{bad.size} mismatches, first at {bad[:5]}"


def test_empty_text(ctx):
    assert ctx.suffix_array_dna(np.zeros(0, np.uint8)).tolist() == [0]  # kiss1_core.hpp:237-238


def test_prefix_doubling_is_the_true_suffix_array(ctx):
    """-s PREFIX_DOUBLING -k -1 is the unique suffix array; with a bounded k it is too when the repeats are shorter than
    what 2^ceil(log2 k) words cover (on longer exact copies the reference's own kISS-2 at k = 256 is not the true array
    either: tests/test_prefix_doubling.py compares with it)"""
    import kiss_b200
    S = tg.planted_repeats(31, 100_000, unit_len=2000, copies=30)
    want = ol.oracle_kiss1(S, k=0xFFFFFFFF)
    assert np.array_equal(ctx.suffix_array_dna(S, 0xFFFFFFFF, kiss_b200.PREFIX_DOUBLING), want)
    S = tg.planted_repeats(31, 100_000, unit_len=300, copies=30)
    assert np.array_equal(ctx.suffix_array_dna(S, 256, kiss_b200.PREFIX_DOUBLING), ol.oracle_kiss1(S, k=0xFFFFFFFF))


def test_ascii_front_end(ctx):
    S = tg.mixed(41, 30_000)
    txt = np.frombuffer(b"ACGT", np.uint8)[S].copy()
    txt[100:110] = ord("N")      # N -> A  (istring.hpp:28-36 gives 4, suffix_sort.hpp:33 takes % 4)
    txt[200:300] |= 0x20         # lower case
    S2 = S.copy()
    S2[100:110] = 0
    want = ol.oracle_kiss1(S2, k=256)
    got = ctx.suffix_array_ascii(txt.tobytes(), 256)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("name,n,gen", [
    ("uniform_5M", 5_000_000, lambda n: tg.uniform(51, n)),
    ("repeats_5M", 5_000_000, lambda n: tg.planted_repeats(52, n, unit_len=5000, copies=100)),
    ("mixed_3M", 3_000_000, lambda n: tg.mixed(53, n)),
    # long homopolymer runs (what a run of N becomes): chains of tens of thousands of waves, written in bursts
    ("long_runs_4M", 4_000_000, lambda n: tg.many_runs(54, n, 60, 5_000, 60_000)),
    ("long_runs_few_2M", 2_000_000, lambda n: tg.many_runs(55, n, 7, 50_000, 200_000)),
], ids=["uniform_5M", "repeats_5M", "mixed_3M", "long_runs_4M", "long_runs_few_2M"])
def test_medium_sizes_bit_exact(ctx, name, n, gen):
    S = gen(n)
    want = ol.oracle_kiss1(S, k=256)
    got = ctx.suffix_array_dna(S, 256)
    bad = np.nonzero(got != want)[0]
    assert bad.size == 0, f"{bad.size} mismatches, first at {bad[:5]}"
    assert ol.fnv1a(got) == ol.fnv1a(want)


@pytest.mark.parametrize("name,gen", [c for c in CASES if c[0] in ("uniform200k", "repeats_mut", "copies1500", "all_T", "uniform7")],
                         ids=["uniform200k", "repeats_mut", "copies1500", "all_T", "uniform7"])
def test_full_sa_into_pinned_buffer(ctx, name, gen):
    """a pinned destination takes the overlapped copy-out: finished SA regions leave on a second stream while
    the S pass is still running"""
    import torch
    S = gen()
    out = torch.empty(S.size + 1, dtype=torch.int32).pin_memory()
    out.fill_(-1)
    sa = ctx.suffix_array_dna(S, 256, out=out.numpy().view(np.uint32))
    assert np.array_equal(sa, ol.oracle_kiss1(S, k=256))


def test_kernel_profile_counts_time_and_algorithmic_bytes(ctx):
    """the per-kernel profiler (bench.py's roofline source): CUDA-event time per kernel class and the launchers'
    own algorithmic byte counts for the radix passes and the tie-group classes; results unchanged when it is on"""
    S = tg.near_identical_copies(19, 150, 1500, 0.005)
    want = ol.oracle_kiss1(S, k=256)
    ctx.profile_enable(True)
    try:
        got = ctx.suffix_array_dna(S, 256)
        prof, alg = ctx.profile(), ctx.profile_alg_bytes()
    finally:
        ctx.profile_enable(False)
    assert np.array_equal(got, want)
    m1 = ctx.last_timings["m"] - 1
    assert prof["k_radix_scatter"][1] >= 4 and prof["k_radix_scatter"][0] > 0
    assert alg["k_radix_scatter"] >= 4 * 16 * (m1 - 400)       # four passes over (key32, pos32), read + write
    assert alg["k_radix_hist"] >= 4 * 4 * (m1 - 400)
    med = {k: v for k, v in alg.items() if k.startswith("k_groups_medium")}
    assert med and all(v % 16 == 0 for v in med.values())     # 16 B per LMS suffix in a refined group
    assert sum(med.values()) >= 16 * 1400                     # the 1500-copy groups went through the block team
    assert set(med) <= set(prof)


def _check_depth_then_position(S, sa, depth, position_rule=True):
    """every neighbour pair of the suffix list `sa` is ordered by the first `depth` bases (a suffix that ends first
    sorts first) and, with position_rule, by position when those are equal; returns the number of such ties"""
    n = S.size
    a, b = sa[:-1].astype(np.int64), sa[1:].astype(np.int64)
    Sp = np.concatenate([S, np.zeros(1, np.uint8)])
    act = np.arange(a.size)
    for j in range(depth):
        if act.size == 0:
            break
        pa, pb = a[act] + j, b[act] + j
        enda, endb = pa >= n, pb >= n
        ca, cb = Sp[np.minimum(pa, n)], Sp[np.minimum(pb, n)]
        wrong = (~enda & endb) | (~enda & ~endb & (ca > cb))
        assert not wrong.any(), f"order violated at depth {j}: list index {act[np.nonzero(wrong)[0][:3]]}"
        decided = enda | (~endb & (ca < cb))
        act = act[~decided]
    if position_rule:
        assert (a[act] < b[act]).all(), "suffixes equal on the whole depth must keep text order"
    return act.size


@pytest.mark.parametrize("k", [32, 256, 300])
def test_prefix_doubling_bounded_k_contract(ctx, k):
    """-s PREFIX_DOUBLING with a bounded k (rank doubling over kISS-2's words, pd.cu): the doubling loop stops at
    offset >= k like the reference's (kiss2_core.hpp:764-766), so LMS suffixes whose first 2^ceil(log2 k) words are equal
    stay tied — in ascending text position here.  The list is a k-order, the induced suffix array is a permutation and a
    k-order, and it is the unique suffix array when no two suffixes share that many bases.  (Equality with the reference's
    own kISS-2 modulo its thread-dependent tie order: tests/test_prefix_doubling.py.)"""
    import kiss_b200
    S = tg.planted_repeats(61, 120_000, unit_len=9000, copies=6)          # exact copies far longer than the words cover
    lms = ctx.sorted_lms(S, k, kiss_b200.PREFIX_DOUBLING)
    assert int(lms[0]) == S.size                                          # the sentinel first (kiss_common.hpp:578)
    assert np.array_equal(np.sort(lms[1:]), ol.oracle_get_lms(S)[0][:-1])
    _check_depth_then_position(S, lms[1:], k, position_rule=False)
    import kiss2_words as kw
    canon, tied_runs = kw.canonical(lms, S, k)                            # runs of equal word classes in text order
    assert tied_runs > 0 and np.array_equal(canon, lms)
    sa = ctx.suffix_array_dna(S, k, kiss_b200.PREFIX_DOUBLING)
    assert int(sa[0]) == S.size and np.array_equal(np.sort(sa), np.arange(S.size + 1, dtype=np.uint32))
    _check_depth_then_position(S, sa, k, position_rule=False)
    S2 = tg.planted_repeats(62, 120_000, unit_len=100, copies=20)         # every LCP below what the words cover
    got = ctx.suffix_array_dna(S2, k, kiss_b200.PREFIX_DOUBLING)
    assert np.array_equal(got, ol.oracle_kiss1(S2, k=0xFFFFFFFF))


@pytest.mark.parametrize("with_ctx", [False, True], ids=["gather_in_waves", "contexts_prefetched"])
@pytest.mark.parametrize("name,gen", [c for c in CASES if c[0] in ("uniform200k", "mixed", "copies1500", "runs300", "uniform7")],
                         ids=["uniform200k", "mixed", "copies1500", "runs300", "uniform7"])
def test_sharded_building_blocks_on_one_gpu(ctx, name, gen, with_ctx):
    """the stage entries of the multi-GPU path as a single shard: prepare -> partition (no splitters) -> subset sort ->
    contexts of the sorted suffixes -> finish from (position, context) pairs = the full suffix array"""
    import torch
    S = gen()
    n = S.size
    dS = torch.zeros(n + 64, dtype=torch.uint8, device="cuda")
    dS[:n] = torch.from_numpy(S).cuda()
    SA = torch.empty(n + 1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    ctx.prepare_dev(dS.data_ptr(), n, 256)
    _, _, n_long = ctx.lms_keys_dev()
    buf = torch.empty(max(n_long, 1), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    counts = ctx.partition_lms_dev(np.zeros(0, np.uint32), 0, n_long, buf.data_ptr())
    assert int(counts[0]) == n_long
    ctx.sort_lms_subset_dev(buf.data_ptr(), n_long)
    cbuf = None
    if with_ctx:
        cbuf = torch.empty(max(n_long, 1), dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        ctx.lms_contexts_dev(buf.data_ptr(), n_long, cbuf.data_ptr())   # enqueued; the next call on the context waits
    ctx.finish_from_sorted_dev(buf.data_ptr(), n_long, SA.data_ptr(), cbuf.data_ptr() if cbuf is not None else None)
    got = SA.cpu().numpy().view(np.uint32)
    assert np.array_equal(got, ol.oracle_kiss1(S, k=256))


@pytest.mark.parametrize("name", ["c1", "c3_200m"])
def test_full_size_golden_checksums_of_the_reference(ctx, name):
    """BASELINE.json configs at sizes the oracle does not finish in seconds: the device suffix array against the digests
    tests/golden/make_full_size_golden.py took from the REFERENCE's own array (48.8 Mbp uniform = the drosophila stand-in,
    200 Mbp chm13-like = the headline text's generator): the device-side position-mixed sum and FNV-1a of the host copy"""
    import json
    import os
    import torch
    g = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "full_size_golden.json")))[name]
    n = g["n"]
    S = tg.uniform_big(g["seed"], n)
    if g["kind"] == "chm13":
        S = tg.chm13_like(g["seed"], n, base=S)
    assert ol.sha_bytes(S[:1 << 20]) == g["text_sha256_head"]          # the same text as the generator script's
    dS = torch.zeros(n + 64, dtype=torch.uint8, device="cuda")
    dS[:n] = torch.from_numpy(S).cuda()
    SA = torch.empty(n + 1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    ctx.suffix_array_dna_dev(dS.data_ptr(), n, SA.data_ptr(), g["k"])
    assert f"{ctx.sa_checksum_dev(SA.data_ptr(), n + 1):016x}" == g["sa_mixed_sum"]
    host = SA.cpu().numpy().view(np.uint32)
    assert f"{ol.fnv1a(host):016x}" == g["sa_fnv"]


def test_sa_checksum_is_the_documented_sum(ctx):
    import torch
    a = tg.sm64(np.arange(100_003, dtype=np.uint64)).astype(np.uint32)
    d = torch.from_numpy(a.view(np.int32)).cuda()
    with np.errstate(over="ignore"):
        want = int(tg.sm64((np.arange(a.size, dtype=np.uint64) << np.uint64(32)) | a.astype(np.uint64)).sum(dtype=np.uint64))
    assert ctx.sa_checksum_dev(d.data_ptr(), a.size) == want
    assert ctx.sa_checksum_dev(d.data_ptr(), 0) == 0


@pytest.mark.parametrize("kname", ["60000", "n-1", "n-2", "n/2", "n/3"])
@pytest.mark.parametrize("name,gen", [c for c in CASES if c[0] in ("uniform200k", "repeats", "mixed", "uniform4097", "uniform100")],
                         ids=["uniform100", "uniform4097", "uniform200k", "repeats", "mixed"])
def test_any_bounded_k_is_accepted(ctx, name, gen, kname):
    """k that is a large part of n: tens of thousands of LMS suffixes have fewer than D(k) bases left and go through the
    reference's depth-shrinking comparator (kiss1_core.hpp:94-135) — a bitonic network on it above 4096 of them; k >= n - 1
    is the unbounded order.  The reference accepts every k; so does this (round 1 refused above 16384 short suffixes)."""
    S = gen()
    n = S.size
    k = {"60000": 60000, "n-1": n - 1, "n-2": n - 2, "n/2": n // 2, "n/3": n // 3}[kname]
    if k < 2:
        pytest.skip("k < 2")
    want = ol.oracle_kiss1(S, k=k)
    got = ctx.suffix_array_dna(S, k)
    bad = np.nonzero(got != want)[0]
    assert bad.size == 0, f"{bad.size} mismatches, first at {bad[:5]}"
    assert np.array_equal(ctx.sorted_lms(S, k), ol.oracle_kiss1(S, k=k, stages=True)["sorted"])
