"""CPU restatement of the reference's FM-index construction and its .fmi file image — TEST INFRASTRUCTURE ONLY (tests/,
never the product path).

Follows FMIndex<SA_INTV = 4, uint32_t, KISS1Sorter<uint32_t>>{.LOOKUP_LEN = 0} as the reference's `fmindex_build` command
instantiates it (include/command/fmindex_build.hpp:27-34):
  build_occ  include/biovoltron/algo/align/exact_match/fm_index.hpp:278-307
  build_bwt  :311-328        build_sa :332-372       build_lookup :237-269 (LOOKUP_LEN = 0 leaves {0, N})
  save       :591-617        Serializer::save include/biovoltron/utility/archive/serializer.hpp:95-112
  DibitVector / XbitVector<1> bit layout include/biovoltron/container/xbit_vector.hpp:21-27, 545-552
The suffix array comes from the caller (the reference builds it with k = 32, fm_index.hpp:385-386).

Pinned: tests/test_fmindex.py compares the image byte for byte with the file the reference itself writes
(oracle/_ref/libkiss_ref.so: ref_fmindex_build_file) and with tests/golden/fmindex_golden.json.
"""
import struct

import numpy as np

OCC1_INTV, OCC2_INTV, B_OCC_INTV = 256, 16, 64


def components(S, SA, sa_intv=4):
    """the members FMIndex::build fills, as numpy arrays"""
    S = np.asarray(S, np.uint8)
    SA = np.asarray(SA, np.uint32)
    N = SA.size                                           # n + 1
    prev = np.where(SA != 0, S[np.maximum(SA.astype(np.int64), 1) - 1], 0).astype(np.uint8)   # :318-324
    pri = int(np.nonzero(SA == 0)[0][0])
    # bwt_: DibitVector<uint8_t>, element e in byte e / 4 at bits 2 (e % 4)
    pad = (-N) % 4
    q = np.concatenate([prev, np.zeros(pad, np.uint8)]).reshape(-1, 4)
    bwt = (q[:, 0] | (q[:, 1] << 2) | (q[:, 2] << 4) | (q[:, 3] << 6)).astype(np.uint8)
    # occ: one-hot counts of the preceding symbol (the entry with SA == 0 counts for nothing, :295-297)
    onehot = np.zeros((N, 4), np.int64)
    onehot[np.arange(N), prev] = 1
    onehot[pri] = 0
    csum = np.concatenate([np.zeros((1, 4), np.int64), np.cumsum(onehot, axis=0)])   # csum[i] = counts over [0, i)
    n1, n2 = N // OCC1_INTV + 1, N // OCC2_INTV + 1
    b1 = np.minimum(np.arange(n1) * OCC1_INTV, N)
    occ1 = csum[b1].astype(np.uint32)                     # exclusive prefix of the block counts (:299-304)
    i2 = np.arange(n2) * OCC2_INTV
    occ2 = (csum[np.minimum(i2, N)] - csum[np.minimum((i2 // OCC1_INTV) * OCC1_INTV, N)]).astype(np.uint8)
    if N % OCC1_INTV == 0:                                # the block loop never reaches beg == N (:283): left at zero
        occ2[N // OCC2_INTV] = 0
    total = csum[N]
    cnt = (1 + np.concatenate([[0], np.cumsum(total)[:-1]])).astype(np.uint32)       # :305-309
    out = dict(cnt=cnt, pri=pri, bwt=bwt, occ1=occ1, occ2=occ2, N=N)
    if sa_intv == 1:
        out["sa"] = SA.copy()
        return out
    hit = (SA % sa_intv) == 0                             # :346-347
    padb = (-N) % 64
    bits = np.concatenate([hit, np.zeros(padb, bool)]).reshape(-1, 64)
    b = (bits.astype(np.uint64) << np.arange(64, dtype=np.uint64)).sum(axis=1, dtype=np.uint64)   # bit e % 64 of word e / 64
    nb = N // B_OCC_INTV + 1
    per = np.zeros(nb, np.int64)
    blocks = bits.sum(axis=1)
    per[:blocks.size] = blocks
    b_occ = (np.cumsum(per) - per).astype(np.uint32)      # :352-356
    sa = np.zeros((N + sa_intv - 1) // sa_intv, np.uint32)   # :358
    vals = SA[hit]
    sa[:vals.size] = vals
    out.update(sa=sa, b=b, b_occ=b_occ)
    return out


def _ser(arr, count):
    """Serializer::save: nothing for an empty range, else the element count as size_t and the raw bytes"""
    if count == 0:
        return b""
    return struct.pack("<Q", count) + np.ascontiguousarray(arr).tobytes()


def fmi_image(S, SA, sa_intv=4):
    """the bytes FMIndex::save writes (fm_index.hpp:591-617) for LOOKUP_LEN = 0"""
    c = components(S, SA, sa_intv)
    N = c["N"]
    img = c["cnt"].astype("<u4").tobytes() + struct.pack("<I", c["pri"])
    img += _ser(c["bwt"], N)
    img += _ser(c["occ1"].astype("<u4"), c["occ1"].shape[0])
    img += _ser(c["occ2"], c["occ2"].shape[0])
    img += _ser(c["sa"].astype("<u4"), c["sa"].size)
    img += _ser(np.array([0, N], "<u4"), 2)               # lookup_ = {0, bwt_.size()} for LOOKUP_LEN = 0
    if sa_intv != 1:
        img += _ser(c["b"].astype("<u8"), N)
        img += _ser(c["b_occ"].astype("<u4"), c["b_occ"].size)
    return img


def layout(N, sa_intv=4):
    """byte offsets of the sections of the image: {name: (offset of the data, bytes)}; every section is preceded by its
    8-byte element count"""
    off = 20
    out = {}
    secs = [("bwt", (N + 3) // 4), ("occ1", 16 * (N // OCC1_INTV + 1)), ("occ2", 4 * (N // OCC2_INTV + 1)),
            ("sa", 4 * (N if sa_intv == 1 else (N + sa_intv - 1) // sa_intv)), ("lookup", 8)]
    if sa_intv != 1:
        secs += [("b", 8 * ((N + 63) // 64)), ("b_occ", 4 * (N // B_OCC_INTV + 1))]
    for name, nbytes in secs:
        out[name] = (off + 8, nbytes)
        off += 8 + nbytes
    out["total"] = off
    return out


def mask_padding(img, N, sa_intv=4):
    """the image with the bits past element N - 1 in the last bwt_ byte and the last b_ word cleared: the reference leaves
    them uninitialised (XbitVector::resize does not clear the tail of its last block; two runs of the reference differ
    there), so parity is defined on everything else"""
    a = bytearray(img)
    L = layout(N, sa_intv)
    assert len(a) == L["total"], (len(a), L["total"])
    if N % 4:
        o, nb = L["bwt"]
        a[o + nb - 1] &= (1 << (2 * (N % 4))) - 1
    if sa_intv != 1 and N % 64:
        o, nb = L["b"]
        w = int.from_bytes(a[o + nb - 8:o + nb], "little") & ((1 << (N % 64)) - 1)
        a[o + nb - 8:o + nb] = w.to_bytes(8, "little")
    return bytes(a)
