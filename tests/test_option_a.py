"""INTEGRATION.md Option A: the reference's own headers and call sites compile UNCHANGED against kiss_b200/host/include
(tools/build_option_a.sh, kiss_b200/host/option_a/option_a_tu.cpp: tests/kiss.cpp:25-28's istring + KISS1Sorter call and
k-order check, KISS2Sorter, and fmindex_build.hpp:27-34's FMIndex<4, uint32_t, KISS1Sorter<uint32_t>>::build + save) and
run on the GPU through the C ABI."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEMO = os.path.join(ROOT, "kiss_b200", "option_a_demo")
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def lcg_text(n=100_000):
    x, out = 12345, np.empty(n, np.uint8)
    for i in range(n):
        x = (x * 1664525 + 1013904223) & 0xFFFFFFFF
        out[i] = (x >> 24) & 3
    return out


@pytest.mark.skipif(not os.path.isdir("/root/reference/include"), reason="the reference tree is not on this machine")
def test_reference_call_sites_compile_against_the_facade():
    subprocess.check_call(["bash", os.path.join(ROOT, "tools", "build_option_a.sh")])
    syms = subprocess.run(["nm", "-D", DEMO], capture_output=True, text=True).stdout
    assert "U kiss_b200_suffix_sort_dna_u32" in syms          # the sorter facades resolved to the C ABI ...


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(DEMO), reason="option_a_demo was not built (needs the reference tree at build time)")
def test_reference_fmindex_over_the_b200_sorter(tmp_path):
    import fmindex_oracle as fo
    import oracle_lib as ol
    out = tmp_path / "a.fmi"
    r = subprocess.run([DEMO, str(out)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.startswith("ok 100001 100001"), r.stdout + r.stderr
    S = lcg_text()
    # the reference's FMIndex code, fed by the GPU suffix array (k = 32), wrote the same file the oracle predicts
    assert fo.mask_padding(out.read_bytes(), S.size + 1) == fo.fmi_image(S, ol.oracle_kiss1(S, k=32))
