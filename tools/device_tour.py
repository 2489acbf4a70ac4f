#!/usr/bin/env python
"""A small tour of every device path (DNA, unbounded, PREFIX_DOUBLING, stage entries, file ingest, FM-index image, general
alphabet, several contexts through the multi entry), each case checked against the CPU oracle; sized so that it also fits
under a memory checker where one is available (compute-sanitizer is closed on this run's GPU pool):
    python tools/device_tour.py [--small]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import general_cases as gc  # noqa: E402
import kiss_b200  # noqa: E402
import oracle_lib as ol  # noqa: E402
import textgen as tg  # noqa: E402

small = "--small" in sys.argv          # racecheck: an order of magnitude slower again
N = 30_000 if small else 120_000
ctx = kiss_b200.Context(0)
done = []


def check(name, got, want):
    ok = bool((np.asarray(got) == np.asarray(want)).all())
    done.append((name, ok))
    print(name, "ok" if ok else "MISMATCH", flush=True)


# DNA path: uniform, repeat-rich (tiny / medium / big groups), homopolymer runs (wave tail), low entropy
for name, S in [("uniform", tg.uniform(1, N)), ("mixed", tg.mixed(2, N)), ("chm13_like", tg.chm13_like(3, N)),
                ("low_entropy", tg.low_entropy(4, N // 4)), ("planted", tg.planted_repeats(5, N, unit_len=3000, copies=12))]:
    check("dna/" + name, ctx.suffix_array_dna(S, 256), ol.oracle_kiss1(S, 256))
S = tg.planted_repeats(6, N, unit_len=4000, copies=40)
check("dna/k32", ctx.suffix_array_dna(S, 32), ol.oracle_kiss1(S, 32))
# unbounded: pivot rounds under the cap, and the fall-back to rank doubling (long exact duplicates)
check("dna/unbounded", ctx.suffix_array_dna(S, -1), ol.oracle_kiss1(S, 0xFFFFFFFF))
D = np.concatenate([tg.uniform(7, N // 2), tg.uniform(7, N // 2), tg.uniform(8, 600)])
check("dna/unbounded_dup", ctx.suffix_array_dna(D, -1), ol.oracle_kiss1(D, 0xFFFFFFFF))
# PREFIX_DOUBLING: unbounded against the oracle; bounded k through its own contract test elsewhere
check("pd/unbounded", ctx.suffix_array_dna(S, -1, kiss_b200.PREFIX_DOUBLING), ol.oracle_kiss1(S, 0xFFFFFFFF))
U = tg.uniform(9, N)
check("pd/k256_uniform", ctx.suffix_array_dna(U, 256, kiss_b200.PREFIX_DOUBLING), ol.oracle_kiss1(U, 256))
ctx.suffix_array_dna(tg.chm13_like(10, N), 64, kiss_b200.PREFIX_DOUBLING)
# stage entries
lms, cnt, lcnt = ctx.get_lms(U)
olms, ocnt, olcnt = ol.oracle_get_lms(U)
check("stage/get_lms", lms, olms)
check("stage/sorted_lms", ctx.sorted_lms(S, 256), ol.oracle_kiss1(S, 256, stages=True)["sorted"])
check("stage/put_lms", ctx.put_lms(S, 256), ol.oracle_kiss1(S, 256, stages=True)["placed"])
# ingest + ASCII + file
import ingest_oracle as io_  # noqa: E402
fa = b">h1 x\n" + bytes(b"ACGT"[c] for c in tg.uniform(11, N // 2)) + b"\nNNNN\n>h2\n" + bytes(b"acgt"[c] for c in tg.uniform(12, 5000)) + b"\n"
sa, n = ctx.suffix_array_file(fa, 256)
codes = io_.read_sequence(fa)
check("file/fasta", sa, ol.oracle_kiss1(codes, 256))
# FM-index image
import fmindex_oracle as fo  # noqa: E402
F = tg.mixed(13, N // 2)
img = ctx.fmindex_build(F, 32)
check("fmindex/image", np.frombuffer(fo.mask_padding(bytes(img), F.size + 1), np.uint8),
      np.frombuffer(fo.mask_padding(fo.fmi_image(F, ol.oracle_kiss1(F, 32)), F.size + 1), np.uint8))
# general alphabet: b = 2, 3, 4
for sig in (5, 20, 200):
    G = gc.hashed(14 + sig, N // 2, np.arange(30, 30 + sig))
    check("general/sigma%d" % sig, ctx.suffix_array(G, 256), ol.oracle_general(G, 256))
# several contexts on one device through the multi entry
ctxs = [kiss_b200.Context(0) for _ in range(3)]
M = tg.chm13_like(15, max(N, 70_000))
check("multi/3ctx", kiss_b200.suffix_array_dna_multi(ctxs, M, 256)[0], ol.oracle_kiss1(M, 256))
bad = [n for n, ok in done if not ok]
print(len(done), "cases,", len(bad), "mismatches", bad)
sys.exit(1 if bad else 0)
