"""kiss_b200_suffix_sort_dna_u32_multi: the sharded LMS sort behind the C ABI (one process, one host thread and one context
per GPU; csrc/multi.cu) against the oracle.  The sharding logic — splitters, stable partition, personalised exchange of
positions, per-range sort, contexts, concatenation on rank 0, induction — is the same whether the contexts sit on
different GPUs or on ONE device, so it is exercised with 2, 3 and 5 contexts on device 0 on a one-GPU box, and across
real GPUs where there are several."""
import subprocess

import numpy as np
import pytest

import oracle_lib as ol
import textgen as tg

pytestmark = pytest.mark.gpu

TEXTS = [
    ("uniform200k", lambda: tg.uniform(7, 200_000)),
    ("repeats", lambda: tg.planted_repeats(8, 300_000, unit_len=3000, copies=40)),
    ("mixed", lambda: tg.mixed(13, 150_000)),
    ("copies5000", lambda: tg.near_identical_copies(20, 60, 5000, 0.003)),      # one huge tie group: a single heavy key range
    ("runs3000", lambda: tg.many_runs(24, 600_000, 3000, 5, 120)),
    ("low_entropy", lambda: np.concatenate([tg.low_entropy(11, 90_000), tg.uniform(12, 1024)])),
]


@pytest.mark.parametrize("ranks", [2, 3, 5])
@pytest.mark.parametrize("name,gen", TEXTS, ids=[t[0] for t in TEXTS])
def test_sharded_sort_with_contexts_on_one_device(name, gen, ranks):
    import kiss_b200
    S = gen()
    ctxs = [kiss_b200.Context(0) for _ in range(ranks)]
    try:
        for k in (256, 32):
            sa, t = kiss_b200.suffix_array_dna_multi(ctxs, S, k)
            want = ol.oracle_kiss1(S, k=k)
            bad = np.nonzero(sa != want)[0]
            assert bad.size == 0, f"k={k}: {bad.size} mismatches, first at {bad[:5]}"
            assert t["m"] == ol.oracle_get_lms(S)[0].size
    finally:
        for c in ctxs:
            c.close()


def test_multi_entry_falls_back_for_global_orders_and_small_texts():
    import kiss_b200
    ctxs = [kiss_b200.Context(0) for _ in range(2)]
    try:
        S = tg.planted_repeats(8, 300_000, unit_len=3000, copies=40)
        sa, _ = kiss_b200.suffix_array_dna_multi(ctxs, S, 0xFFFFFFFF)
        assert np.array_equal(sa, ol.oracle_kiss1(S, k=0xFFFFFFFF))
        sa, _ = kiss_b200.suffix_array_dna_multi(ctxs, S, 256, kiss_b200.PREFIX_DOUBLING)
        assert np.array_equal(sa, ctxs[0].suffix_array_dna(S, 256, kiss_b200.PREFIX_DOUBLING))
        S = tg.uniform(3, 1000)
        sa, _ = kiss_b200.suffix_array_dna_multi(ctxs, S, 256)
        assert np.array_equal(sa, ol.oracle_kiss1(S, k=256))
        sa, _ = kiss_b200.suffix_array_dna_multi(ctxs, np.zeros(0, np.uint8), 256)
        assert sa.tolist() == [0]
    finally:
        for c in ctxs:
            c.close()


def test_sharded_sort_across_real_gpus():
    import torch
    import kiss_b200
    g = torch.cuda.device_count()
    if g < 2:
        pytest.skip("needs at least 2 GPUs")
    ctxs = [kiss_b200.Context(d) for d in range(g)]
    try:
        S = tg.chm13_like(3, 20_000_000)
        sa, t = kiss_b200.suffix_array_dna_multi(ctxs, S, 256)
        assert np.array_equal(sa, ol.oracle_kiss1(S, k=256))
    finally:
        for c in ctxs:
            c.close()


def test_cli_gpus_flag(tmp_path):
    """`kISS --gpus N suffix_sort` (an opt-in flag that does not change the defaults)"""
    import os
    import torch
    from test_cli import KISS
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    S = tg.mixed(61, 200_000)
    f = tmp_path / "t.txt"
    f.write_bytes(np.frombuffer(b"ACGT", np.uint8)[S].tobytes() + b"\n")
    out = tmp_path / "sa.bin"
    r = subprocess.run([KISS, "--gpus", "2", "suffix_sort", str(f), "--out", str(out)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert np.array_equal(np.fromfile(out, dtype=np.uint32), ol.oracle_kiss1(S, k=256))
