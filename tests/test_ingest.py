"""The ingest step: file bytes -> codes with the reference's read_sequence semantics.
CPU: the restatement (oracle/ingest_oracle.py) against the reference itself and the committed golden vectors.
GPU (`-m gpu`): the device kernels through the C ABI against the restatement, and the whole file -> suffix array call."""
import json
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE), "oracle"))
import ingest_cases as ic  # noqa: E402
import ingest_oracle as io_  # noqa: E402
import oracle_lib as ol  # noqa: E402

ALL = dict(ic.CASES)
ALL.update(ic.big_cases())
GOLDEN = json.load(open(os.path.join(HERE, "golden", "ingest_golden.json")))


@pytest.mark.parametrize("name", sorted(ALL))
def test_oracle_matches_golden_vectors_from_the_reference(name):
    codes = io_.read_sequence(ALL[name])
    g = GOLDEN[name]
    assert g["bytes"] == len(ALL[name])
    assert codes.size == g["n"] and codes[:32].tolist() == g["head"] and ol.sha_bytes(codes) == g["sha256"]


@pytest.mark.skipif(not ol.have_ref() or not hasattr(ol.ref(), "ref_read_sequence"), reason="reference harness not built")
@pytest.mark.parametrize("name", sorted(ALL))
def test_oracle_matches_reference_read_sequence(name):
    assert np.array_equal(io_.read_sequence(ALL[name]), ol.ref_read_sequence(ALL[name]))


@pytest.fixture(scope="module")
def ctx():
    import kiss_b200
    c = kiss_b200.Context(0)
    yield c
    c.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(ALL))
def test_device_ingest_matches_oracle(ctx, name):
    import torch
    data = ALL[name]
    want = io_.read_sequence(data)
    raw = torch.zeros(len(data) + 16, dtype=torch.uint8, device="cuda")
    if data:
        raw[:len(data)] = torch.frombuffer(bytearray(data), dtype=torch.uint8).cuda()
    codes = torch.full((len(data) + 16,), 77, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    n = ctx.ingest_dev(raw.data_ptr(), len(data), data[:1] == b">", codes.data_ptr())
    got = codes.cpu().numpy()
    assert n == want.size
    assert np.array_equal(got[:n], want)
    assert (got[n:] == 77).all()          # nothing written past the text


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["one_record", "crlf", "header_only", "empty", "text_mode", "fasta_60", "fasta_lower_n",
                                  "fasta_wide_lines", "long_header"])
def test_file_to_suffix_array(ctx, name):
    data = ALL[name]
    S = io_.read_sequence(data)
    sa, n = ctx.suffix_array_file(data, 256)
    assert n == S.size
    assert np.array_equal(sa, ol.oracle_kiss1(S, k=256))
