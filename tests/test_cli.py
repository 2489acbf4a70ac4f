"""`kISS suffix_sort` command (kiss_b200/host/kiss_main.cpp) against the reference CLI's semantics
(include/utils/options.hpp, include/command/suffix_sort.hpp, include/utils/io.hpp)."""
import os
import re
import subprocess

import numpy as np
import pytest

import oracle_lib as ol
import textgen as tg

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KISS = os.path.join(ROOT, "kiss_b200", "kISS")


def run(*args):
    return subprocess.run([KISS, *args], capture_output=True, text=True, timeout=300)


def test_version_and_help_exit_1():
    r = run("-v")
    assert r.returncode == 1 and r.stdout.strip() == "1.0.0"     # options.hpp:255-264
    r = run("--help")
    assert r.returncode == 1 and "suffix_sort" in r.stdout
    r = run()
    assert r.returncode == 1


def test_bad_arguments_print_help_and_exit_1():
    r = run("suffix_sort", "-s", "bogus", "x.fa")                # options.hpp:67-81, 240-244
    assert r.returncode == 1 and "sorting-algorithm" in r.stderr
    r = run("suffix_sort", "--nope", "x.fa")
    assert r.returncode == 1


def test_generic_flag_is_refused(tmp_path):
    f = tmp_path / "t.txt"
    f.write_text("ACGT\n")
    r = run("-g", "suffix_sort", str(f))                         # suffix_sort.hpp:26-28
    assert r.returncode != 0 and "Generic sorting and indexing are currently not supported." in r.stderr


def test_unknown_command():
    r = run("frobnicate", "x")
    assert r.returncode != 0 and "frobnicate" in r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["fasta", "text"])
def test_suffix_sort_file(tmp_path, mode):
    S = tg.mixed(61, 50_000)
    txt = np.frombuffer(b"ACGT", np.uint8)[S].copy()
    txt[500:520] = ord("N")                                      # N -> A
    txt[1000:1100] |= 0x20                                       # lower case
    S[500:520] = 0
    lines = [txt[i:i + 70].tobytes().decode() for i in range(0, txt.size, 70)]
    f = tmp_path / ("t.fa" if mode == "fasta" else "t.txt")
    if mode == "fasta":                                          # io.hpp:8-11: records concatenated, headers dropped
        half = len(lines) // 2
        f.write_text(">chr1 first\n" + "\n".join(lines[:half]) + "\n>chr2\n" + "\n".join(lines[half:]) + "\n")
    else:
        f.write_text("\n".join(lines) + "\n")
    out = tmp_path / "sa.bin"
    for algo, k in (("parallel_sorting", "256"), ("PREFIX_DOUBLING", "-1")):
        r = run("--verbose", "suffix_sort", str(f), "-k", k, "-s", algo, "--out", str(out))
        assert r.returncode == 0, r.stderr
        kk = 256 if k == "256" else 2 ** 64 - 1
        assert re.search(rf"\[info\] \[suffix_sort.hpp:61\] n = {S.size}, k = {kk}, suffix sorting elapsed [0-9.]+", r.stderr)
        assert "induced_sort elapsed" in r.stderr
        sa = np.fromfile(out, dtype=np.uint32)
        assert np.array_equal(sa, ol.oracle_kiss1(S, k=0xFFFFFFFF if k == "-1" else 256))


@pytest.mark.gpu
def test_empty_file(tmp_path):
    f = tmp_path / "e.txt"
    f.write_text("")
    out = tmp_path / "sa.bin"
    r = run("suffix_sort", str(f), "--out", str(out))
    assert r.returncode == 0 and "n = 0" in r.stderr
    assert np.fromfile(out, dtype=np.uint32).tolist() == [0]     # kiss1_core.hpp:237-238


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["fasta_80_crlf", "blank_lines", "gt_inside_line", "long_header", "text_mode_gt_line",
                                  "two_headers_in_a_row", "header_run_across_chunks", "long_header_run_odd"])
def test_device_ingest_equals_host_ingest(tmp_path, name):
    """default = the file's bytes are stripped / encoded on the device (read_sequence semantics, io.hpp:6-18);
    --host-ingest = the serial getline loop on the host: same n, same suffix array"""
    import ingest_cases as ic
    data = dict(ic.CASES, **ic.big_cases())[name]
    f = tmp_path / "x.fa"
    f.write_bytes(data)
    outs = []
    for extra in ([], ["--host-ingest"]):
        out = tmp_path / ("sa%d.bin" % len(outs))
        r = run("suffix_sort", str(f), "--out", str(out), *extra)
        assert r.returncode == 0, r.stderr
        n = int(re.search(r"n = (\d+),", r.stderr).group(1))
        outs.append((n, np.fromfile(out, dtype=np.uint32)))
    assert outs[0][0] == outs[1][0] and np.array_equal(outs[0][1], outs[1][1])
    assert outs[0][1].size == outs[0][0] + 1


@pytest.mark.gpu
def test_fmindex_build_writes_the_reference_fmi_file(tmp_path):
    """`kISS fmindex_build <file>` (include/command/fmindex_build.hpp:14-36) writes <file>.fmi: the bytes of the reference's
    FMIndex<4, uint32_t, KISS1Sorter<uint32_t>>{.LOOKUP_LEN = 0}::save — compared with the file the reference itself writes
    (oracle/_ref) when the harness is there, and with the restatement"""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import fmindex_oracle as fo
    import ingest_cases as ic
    import ingest_oracle as io_
    data = ic.big_cases()["fasta_60"]
    f = tmp_path / "g.fa"
    f.write_bytes(data)
    r = run("--verbose", "fmindex_build", str(f))
    assert r.returncode == 0, r.stderr
    S = io_.read_sequence(data)
    assert re.search(rf"n = {S.size}, \d+ bytes, index building elapsed", r.stderr)
    img = (tmp_path / "g.fa.fmi").read_bytes()
    assert img == fo.fmi_image(S, ol.oracle_kiss1(S, k=32))
    if ol.have_ref() and hasattr(ol.ref(), "ref_fmindex_build_file"):
        assert img == fo.mask_padding(ol.ref_fmindex_image(S), S.size + 1)


@pytest.mark.gpu
@pytest.mark.parametrize("algo", ["PARALLEL_SORTING", "PREFIX_DOUBLING"])
def test_general_extension_sorts_a_protein_fasta(tmp_path, algo):
    """--general (an extension: the reference's CLI refuses -g): the sequence lines' bytes as they are, through the
    general-alphabet sorters (kiss1_sorter.hpp:28-44); records concatenated and headers dropped like read_sequence"""
    import general_cases as gc
    S = gc.hashed(71, 40_000, gc.AMINO)
    lines = [S[i:i + 60].tobytes().decode() for i in range(0, S.size, 60)]
    half = len(lines) // 2
    f = tmp_path / "p.fa"
    f.write_text(">sp|P1 first\n" + "\n".join(lines[:half]) + "\n>sp|P2 second\n" + "\n".join(lines[half:]) + "\n")
    out = tmp_path / "sa.bin"
    r = run("suffix_sort", str(f), "--general", "-k", "256", "-s", algo, "--out", str(out))
    assert r.returncode == 0, r.stderr
    assert f"n = {S.size}" in r.stderr
    sa = np.fromfile(out, dtype=np.uint32)
    assert np.array_equal(sa, ol.oracle_general(S, 256))
