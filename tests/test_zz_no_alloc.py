"""kiss_b200_ingest_noalloc_dev: the allocation-free variant of the device ingest (no counterpart in the reference).
Its scratch is a fixed __device__ array of the library, so it is compared with the
existing entry kiss_b200_ingest_dev (arena scratch) — byte for byte, the path is integer work — on two calls in a row (the
array keeps the previous call's contents), and sizes beyond the stated maximum must be refused.

Named test_zz_* so that pytest collects it last."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE), "oracle"))
import ingest_cases as ic  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import kiss_b200
    c = kiss_b200.Context(0)
    yield c
    c.close()


def _both(ctx, data):
    """(n, codes) from the existing entry and from the allocation-free one"""
    import torch
    raw = torch.zeros(len(data) + 16, dtype=torch.uint8, device="cuda")
    if data:
        raw[:len(data)] = torch.frombuffer(bytearray(data), dtype=torch.uint8).cuda()
    fasta = data[:1] == b">"
    out = []
    for fn in (ctx.ingest_dev, ctx.ingest_noalloc_dev):
        codes = torch.full((len(data) + 16,), 77, dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        n = fn(raw.data_ptr(), len(data), fasta, codes.data_ptr())
        out.append((n, codes.cpu().numpy()))
    return out


def test_noalloc_ingest_equals_ingest_dev_on_two_calls_in_a_row(ctx):
    big = ic.big_cases()
    # a large multi-record file, then a smaller one with header runs (stale scratch of the first call must not matter),
    # then the first again
    seq = [big["fasta_60"], ic.CASES["header_run_across_chunks"], big["fasta_60"], big["fasta_lower_n"],
           ic.CASES["one_record"], big["text_big"]]
    for data in seq:
        (n0, c0), (n1, c1) = _both(ctx, data)
        assert n1 == n0
        assert np.array_equal(c1, c0)          # bit-exact, including the untouched tail (77)


@pytest.mark.parametrize("name", ["empty", "header_only", "two_headers_in_a_row", "long_header_run_odd", "text_mode", "crlf"])
def test_noalloc_ingest_small_cases(ctx, name):
    (n0, c0), (n1, c1) = _both(ctx, ic.CASES[name])
    assert n1 == n0 and np.array_equal(c1, c0)


def test_noalloc_ingest_refuses_more_than_its_fixed_scratch(ctx):
    import kiss_b200
    from kiss_b200 import _cabi
    with pytest.raises(kiss_b200.KissB200Error) as e:
        ctx.ingest_noalloc_dev(256, _cabi.NOALLOC_MAX_BYTES + 1, True, 256)   # refused before any pointer is used
    assert e.value.code == -2 and "KISS_B200_NOALLOC_MAX_BYTES" in str(e.value)
