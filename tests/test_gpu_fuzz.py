"""Seeded fuzz of the whole path through the C ABI against the CPU oracle: small texts of every shape (n from 0, skewed
alphabets, runs, periodic pieces, planted copies), k from 1 to beyond n and -1, both algorithms where their results are
defined to coincide (the true suffix array).  Complements the named cases of test_gpu_parity.py with the ragged sizes
nobody thought of."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import general_cases as gc
import oracle_lib as ol
import textgen as tg

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import kiss_b200
    return kiss_b200.Context(0)


def fuzz_text(r, sigma=4):
    """a text assembled from 1-6 pieces: uniform, skewed, one-symbol run, periodic, a copy of an earlier stretch"""
    n_target = [0, 1, 2, 3, 5, 17, 64, 200, 1000, 3000, 9000][r.below(11)] + r.below(7)
    parts, total = [], 0
    while total < n_target:
        kind = r.below(5)
        ln = min(n_target - total, 1 + r.below(max(1, n_target // 2)))
        if kind == 0:
            p = gc.hashed(r.u64() & 0xFFFF, ln, np.arange(sigma))
        elif kind == 1:
            p = (gc.hashed(r.u64() & 0xFFFF, ln, np.arange(4)) // 3 * r.below(sigma)).astype(np.uint8)   # mostly one symbol
        elif kind == 2:
            p = np.full(ln, r.below(sigma), np.uint8)
        elif kind == 3:
            unit = np.array([r.below(sigma) for _ in range(1 + r.below(5))], np.uint8)
            p = np.tile(unit, ln // unit.size + 1)[:ln]
        else:
            if not parts:
                continue
            src = np.concatenate(parts)
            a = r.below(src.size)
            p = src[a:a + ln].copy()
            if p.size == 0:
                continue
        parts.append(p.astype(np.uint8))
        total += p.size
    return np.concatenate(parts).astype(np.uint8) if parts else np.zeros(0, np.uint8)


def fuzz_k(r, n):
    c = r.below(6)
    if c == 0:
        return 0xFFFFFFFF
    if c == 1:
        return 1 + r.below(8)
    if c == 2:
        return 120 + r.below(12)          # around the 125-base block boundary
    if c == 3:
        return max(1, n - 2 + r.below(5))  # around n
    if c == 4:
        return 256
    return 1 + r.below(max(1, 2 * n))


@pytest.mark.parametrize("seed", range(8))
def test_dna_path_against_the_oracle(ctx, seed):
    import kiss_b200
    r = tg._Rng(1000 + seed)
    for _ in range(40):
        S = fuzz_text(r)
        k = fuzz_k(r, S.size)
        want = ol.oracle_kiss1(S, k)
        got = ctx.suffix_array_dna(S, k)
        assert np.array_equal(got, want), f"seed {seed}: n={S.size} k={k} text={S[:80].tolist()}"
        if k == 0xFFFFFFFF or k + 1 >= S.size:   # the true suffix array: PREFIX_DOUBLING must give the same
            got2 = ctx.suffix_array_dna(S, k, kiss_b200.PREFIX_DOUBLING)
            assert np.array_equal(got2, want), f"seed {seed} (PREFIX_DOUBLING): n={S.size} k={k} text={S[:80].tolist()}"


@pytest.mark.parametrize("seed", range(4))
def test_general_path_full_order_against_the_oracle(ctx, seed):
    r = tg._Rng(2000 + seed)
    for _ in range(25):
        sigma = [1, 2, 3, 4, 5, 16, 17, 64, 65, 200][r.below(10)]
        S = fuzz_text(r, sigma)
        S = (S.astype(np.uint16) * (255 // max(1, sigma - 1) if sigma > 1 else 1)).astype(np.uint8)   # spread over the byte range
        want = ol.oracle_general(S, gc.FULL)
        for algo in (0, 1):
            got = ctx.suffix_array(S, gc.FULL, algo)
            assert np.array_equal(got, want), f"seed {seed}: n={S.size} sigma={sigma} algo={algo} text={S[:60].tolist()}"


@pytest.mark.parametrize("seed", range(4))
def test_fmindex_image_against_the_oracle(ctx, seed):
    import fmindex_oracle as fo
    r = tg._Rng(3000 + seed)
    for _ in range(12):
        S = fuzz_text(r)
        if S.size == 0:
            continue
        intv = [1, 2, 3, 4, 5, 8, 16][r.below(7)]
        img = ctx.fmindex_build(S, 32, sa_intv=intv)
        want = fo.fmi_image(S, ol.oracle_kiss1(S, 32), intv)
        N = S.size + 1
        assert fo.mask_padding(bytes(img), N, intv) == fo.mask_padding(want, N, intv), f"seed {seed}: n={S.size} intv={intv}"


@pytest.mark.parametrize("seed", range(6))
def test_device_ingest_against_the_oracle(ctx, seed):
    """files made of random lines — headers, header runs, empty lines, CR LF, missing final newline, '>' inside lines —
    through kiss_b200_suffix_sort_file_u32: the text the device extracts must be read_sequence's (io.hpp:6-18)"""
    import ingest_oracle as io_
    r = tg._Rng(4000 + seed)
    alphabet = b"ACGTacgtNn>\r xyz"
    for _ in range(30):
        lines = []
        for _ in range(r.below(12)):
            c = r.below(10)
            if c < 2:
                lines.append(b">" + bytes(alphabet[r.below(len(alphabet))] for _ in range(r.below(6))))
            elif c == 2:
                lines.append(b"")
            else:
                ln = [0, 1, 2, 5, 60, 61, 300, 4096, 5000][r.below(9)]
                lines.append(bytes(alphabet[r.below(9 if r.below(4) else len(alphabet))] for _ in range(ln)))
        data = (b"\r\n" if r.below(5) == 0 else b"\n").join(lines)
        if r.below(3):
            data += b"\n"
        want_text = io_.read_sequence(data)
        sa, n = ctx.suffix_array_file(data, 256)
        assert n == want_text.size, f"seed {seed}: {data[:120]!r}"
        assert np.array_equal(sa, ol.oracle_kiss1(want_text, 256)), f"seed {seed}: {data[:120]!r}"


@pytest.mark.skipif(not ol.have_ref() or not hasattr(ol.ref(), "ref_kiss2_sorted_lms"), reason="reference harness not built")
@pytest.mark.parametrize("seed", range(4))
def test_bounded_prefix_doubling_against_kiss2_itself(ctx, seed):
    """bounded-k PREFIX_DOUBLING: the LMS order equals the reference's kISS-2 (single-threaded, its deterministic run) after
    putting every run of equal word classes in text order (tests/kiss2_words.py)"""
    import kiss2_words as kw
    import kiss_b200
    r = tg._Rng(5000 + seed)
    done = 0
    while done < 25:
        S = fuzz_text(r)
        if S.size < 64:          # (below 64 bases the entry takes the pivot path: the reference's tail-word rule degenerates)
            continue
        k = [1, 2, 3, 8, 16, 17, 32, 100, 256][r.below(9)]
        ref_lms = ol.ref_kiss2_sorted_lms(S, k, 1)
        got_lms = ctx.sorted_lms(S, k, kiss_b200.PREFIX_DOUBLING)
        want, _ = kw.canonical(ref_lms, S, k)
        assert np.array_equal(got_lms, want), f"seed {seed}: n={S.size} k={k} text={S[:80].tolist()}"
        done += 1
