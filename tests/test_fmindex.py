"""FM-index construction: the .fmi image of the reference's `fmindex_build` command.
CPU: the restatement (oracle/fmindex_oracle.py) byte for byte against the file the REFERENCE ITSELF writes (oracle/_ref:
FMIndex<4, uint32_t, KISS1Sorter<uint32_t>>{.LOOKUP_LEN = 0}.build + save) and against tests/golden/fmindex_golden.json.
GPU (`-m gpu`): the device construction through the C ABI against the restatement.
The reference leaves the bits past the last element of bwt_ / b_ uninitialised (two of its own runs differ there):
mask_padding() clears them before a reference image is compared."""
import hashlib
import json
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE), "oracle"))
import fmindex_oracle as fo  # noqa: E402
import oracle_lib as ol  # noqa: E402
import textgen as tg  # noqa: E402

CASES = {
    "uniform1": lambda: tg.uniform(1, 1),
    "uniform2": lambda: tg.uniform(2, 2),
    "uniform3": lambda: tg.uniform(3, 3),
    "uniform15": lambda: tg.uniform(4, 15),           # N = 16: one occ2 row boundary exactly
    "uniform63": lambda: tg.uniform(9, 63),           # N = 64: one b_occ block exactly
    "uniform255": lambda: tg.uniform(5, 255),         # N = 256: the occ1 row loop never reaches the last row
    "uniform256": lambda: tg.uniform(5, 256),
    "uniform1000": lambda: tg.uniform(5, 1000),
    "uniform4097": lambda: tg.uniform(6, 4097),
    "uniform200k": lambda: tg.uniform(7, 200_000),
    "all_A": lambda: np.zeros(500, np.uint8),
    "all_T": lambda: np.full(511, 3, np.uint8),
    "repeats": lambda: tg.planted_repeats(8, 300_000, unit_len=3000, copies=40),   # ties deeper than k = 32
    "tandem": lambda: tg.tandem(10, 200_000),
    "mixed": lambda: tg.mixed(13, 150_000),
}
NAMES = sorted(CASES)
GOLDEN_PATH = os.path.join(HERE, "golden", "fmindex_golden.json")
needs_ref = pytest.mark.skipif(not ol.have_ref() or not hasattr(ol.ref(), "ref_fmindex_build_file"),
                               reason="reference harness not built")


def oracle_image(S, sa_intv=4):
    return fo.fmi_image(S, ol.oracle_kiss1(S, k=32), sa_intv)      # k = 32: fm_index.hpp:385-386


@needs_ref
@pytest.mark.parametrize("name", NAMES)
def test_oracle_image_equals_the_reference_file(name):
    S = CASES[name]()
    ref = ol.ref_fmindex_image(S)
    assert fo.mask_padding(ref, S.size + 1) == oracle_image(S)


@pytest.mark.parametrize("name", NAMES)
def test_oracle_image_matches_golden(name):
    g = json.load(open(GOLDEN_PATH))[name]
    img = oracle_image(CASES[name]())
    assert len(img) == g["bytes"] and hashlib.sha256(img).hexdigest() == g["sha256"]


@pytest.fixture(scope="module")
def ctx():
    import kiss_b200
    c = kiss_b200.Context(0)
    yield c
    c.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_device_fmindex_image_equals_oracle(ctx, name):
    S = CASES[name]()
    got = ctx.fmindex_build(S).tobytes()
    want = oracle_image(S)
    assert len(got) == len(want)
    if got != want:
        L = fo.layout(S.size + 1)
        bad = next(i for i in range(len(got)) if got[i] != want[i])
        sec = [k for k, v in L.items() if k != "total" and v[0] - 8 <= bad < v[0] + v[1]]
        pytest.fail(f"first difference at byte {bad} (section {sec or 'header'})")


@pytest.mark.gpu
@pytest.mark.parametrize("sa_intv", [1, 2, 3, 8, 10, 32])
@pytest.mark.parametrize("name", ["uniform1", "uniform255", "uniform4097", "repeats"])
def test_device_fmindex_other_sampling_intervals(ctx, name, sa_intv):
    S = CASES[name]()
    assert ctx.fmindex_build(S, sa_intv=sa_intv).tobytes() == oracle_image(S, sa_intv)


@pytest.mark.gpu
def test_fmindex_from_file_and_from_device_sa(ctx):
    import torch
    import ingest_cases as ic
    import ingest_oracle as io_
    data = ic.big_cases()["fasta_lower_n"]
    S = io_.read_sequence(data)
    img, n = ctx.fmindex_build_file(data)
    assert n == S.size and img.tobytes() == oracle_image(S)
    # any suffix array given on the device (here: the k = 256 one)
    sa = ol.oracle_kiss1(S, k=256)
    dS = torch.zeros(S.size + 64, dtype=torch.uint8, device="cuda")
    dS[:S.size] = torch.from_numpy(S).cuda()
    dSA = torch.from_numpy(sa.view(np.int32)).cuda()
    torch.cuda.synchronize()
    assert ctx.fmindex_from_sa_dev(dS.data_ptr(), S.size, dSA.data_ptr()).tobytes() == fo.fmi_image(S, sa)


@pytest.mark.gpu
def test_fmindex_image_buffer_too_small_is_refused(ctx):
    import kiss_b200
    S = tg.uniform(7, 1000)
    with pytest.raises(kiss_b200.KissB200Error):
        ctx.fmindex_build(S, out=np.empty(100, np.uint8))
