"""Golden vectors generated from the reference itself (tests/golden/make_golden.py): the oracle on the CPU,
the CUDA path on the GPU."""
import json
import os
import re

import numpy as np
import pytest

import oracle_lib as ol
import textgen as tg

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLD = json.load(open(os.path.join(HERE, "golden", "kiss1_golden.json")))["vectors"]
IDS = [f"{v['gen']}{v['n']}-k{v['k']}" for v in GOLD]


def _text(v):
    S = getattr(tg, v["gen"])(**v["args"])
    if v["gen"] == "low_entropy":
        S = np.concatenate([S, tg.uniform(v["args"]["seed"] + 1, 1024)])
    assert S.size == v["n"]
    return S


@pytest.mark.parametrize("v", GOLD, ids=IDS)
def test_oracle_matches_reference_golden(v):
    o = ol.oracle_kiss1(_text(v), k=v["k"], stages=True)
    assert o["m"] == v["m"]
    assert o["count"].tolist() == v["count"] and o["lms_count"].tolist() == v["lms_count"]
    assert f"{ol.fnv1a(o['lms']):016x}" == v["lms_fnv"]
    assert f"{ol.fnv1a(o['sorted']):016x}" == v["sorted_fnv"]
    assert f"{ol.fnv1a(o['sa']):016x}" == v["sa_fnv"]
    assert o["sa"][:8].tolist() == v["sa_head"] and o["sa"][-8:].tolist() == v["sa_tail"]


@pytest.mark.gpu
@pytest.mark.parametrize("v", GOLD, ids=IDS)
def test_cuda_path_matches_reference_golden(v):
    import kiss_b200
    ctx = kiss_b200.default_context()
    S = _text(v)
    sa = ctx.suffix_array_dna(S, v["k"])
    assert f"{ol.fnv1a(sa):016x}" == v["sa_fnv"]
    assert sa[:8].tolist() == v["sa_head"] and sa[-8:].tolist() == v["sa_tail"]
    srt = ctx.sorted_lms(S, v["k"])
    assert srt.size == v["m"] and f"{ol.fnv1a(srt):016x}" == v["sorted_fnv"]


def test_library_exports_every_declared_symbol():
    """the C-ABI library loads without a GPU and exports exactly what include/kiss_b200.h declares"""
    import kiss_b200
    from kiss_b200 import _cabi
    hdr = open(os.path.join(ROOT, "include", "kiss_b200.h")).read()
    declared = set(re.findall(r"\b(kiss_b200_[a-z0-9_]+)\s*\(", hdr))
    L = kiss_b200.lib()
    for name in sorted(declared):
        assert hasattr(L, name), f"{name} declared in kiss_b200.h but not exported"
    assert declared == set(_cabi.SIGNATURES), declared ^ set(_cabi.SIGNATURES)
    assert L.kiss_b200_abi_version() == 1
    assert L.kiss_b200_strerror(-6).decode().startswith("no CUDA device")


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import kiss_b200
    with pytest.raises(kiss_b200.KissB200Error):
        kiss_b200.Context(0)


def test_product_does_not_touch_the_oracle():
    """only tests/, bench.py's CPU legs and __graft_entry__.smoke() may load anything under oracle/"""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "kiss_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp", "Makefile")):
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                assert "oracle_lib" not in txt and "libkiss_oracle" not in txt and "libkiss_ref" not in txt, f
