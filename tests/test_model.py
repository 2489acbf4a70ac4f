"""The numpy model of the GPU decomposition (tests/model_gpu_algo.py) against the oracle."""
import numpy as np
import pytest

import model_gpu_algo as mg
import oracle_lib as ol
import textgen as tg

CASES = [
    lambda: tg.uniform(1, 1), lambda: tg.uniform(2, 2), lambda: tg.uniform(3, 9),
    lambda: tg.uniform(4, 300), lambda: tg.uniform(5, 3000),
    lambda: tg.planted_repeats(6, 20000, unit_len=1500, copies=8),
    lambda: tg.tandem(7, 20000, arrays=8, max_unit=9, max_len=3000),
    lambda: tg.mixed(8, 12000),
    lambda: np.zeros(300, np.uint8), lambda: np.full(300, 3, np.uint8),
    lambda: np.concatenate([tg.periodic(3000, [1, 0]), tg.uniform(9, 600)]),
    lambda: tg.low_entropy(10, 700, 2),
]


@pytest.mark.parametrize("k", [256, 32, 3, 0xFFFFFFFF])
@pytest.mark.parametrize("ci", range(len(CASES)))
def test_model_equals_oracle(ci, k):
    S = CASES[ci]()
    o = ol.oracle_kiss1(S, k=k, stages=True) if S.size else None
    sa = mg.model_suffix_array(S, k)
    exp = o["sa"] if o else np.zeros(1, np.uint32)
    if o:
        lms = mg.lms_positions(S)
        assert np.array_equal(lms, o["lms"][:-1])
    assert np.array_equal(sa, exp)


def test_uniform_big_is_the_same_stream():
    import textgen as tg
    for seed, n in [(3, 0), (3, 1), (5, 33), (9, 100_001)]:
        assert np.array_equal(tg.uniform_big(seed, n), tg.uniform(seed, n))
    assert np.array_equal(tg.uniform_big(11, 300, step=64), tg.uniform(11, 300))
