"""ctypes bindings for the CPU checkers: oracle/libkiss_oracle.so (plain-C restatement) and,
when present, oracle/_ref/libkiss_ref.so (the unmodified reference compiled by oracle/build_ref.sh).
Test infrastructure only."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
_ORACLE_SO = os.path.join(ORACLE_DIR, "libkiss_oracle.so")
_REF_SO = os.path.join(ORACLE_DIR, "_ref", "libkiss_ref.so")

EMPTY = 0xFFFFFFFF
_vp, _u32, _u64, _dbl = C.c_void_p, C.c_uint32, C.c_uint64, C.c_double


def _ptr(a):
    return None if a is None else a.ctypes.data


def build_oracle():
    srcs = [os.path.join(ORACLE_DIR, f) for f in os.listdir(ORACLE_DIR) if f.endswith("_oracle.c")]
    if not os.path.exists(_ORACLE_SO) or any(
            os.path.getmtime(s) > os.path.getmtime(_ORACLE_SO) for s in srcs):
        subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, os.path.join(ORACLE_DIR, "libkiss_oracle.so")])


_oracle = None
_ref = None


def oracle():
    global _oracle
    if _oracle is None:
        build_oracle()
        L = C.CDLL(_ORACLE_SO)
        L.oracle_get_lms.restype = _u32
        L.oracle_get_lms.argtypes = [_vp, _u32, _vp, _vp, _vp]
        L.oracle_lms_less.restype = C.c_int
        L.oracle_lms_less.argtypes = [_vp, _u32, _u32, _u32, _u32]
        L.oracle_lms_suffix_direct_sort_dna.restype = None
        L.oracle_lms_suffix_direct_sort_dna.argtypes = [_vp, _u32, _vp, _u32, _u32]
        L.oracle_put_lms_suffix.restype = None
        L.oracle_put_lms_suffix.argtypes = [_vp, _u32, _u32, _vp, _vp]
        L.oracle_induced_L.restype = None
        L.oracle_induced_L.argtypes = [_vp, _u32, _vp, _vp]
        L.oracle_induced_clear.restype = None
        L.oracle_induced_clear.argtypes = [_u32, _vp, _vp, _vp]
        L.oracle_induced_S.restype = None
        L.oracle_induced_S.argtypes = [_vp, _u32, _vp, _vp]
        L.oracle_kiss1_suffix_array_dna.restype = _u32
        L.oracle_kiss1_suffix_array_dna.argtypes = [_vp, _u32, _u32, _vp, _vp, _vp, _vp, _vp, _vp]
        L.oracle_induce_from_sorted_lms.restype = None
        L.oracle_induce_from_sorted_lms.argtypes = [_vp, _u32, _vp, _u32, _vp]
        L.oracle_kiss1_suffix_array.restype = _u32
        L.oracle_kiss1_suffix_array.argtypes = [_vp, _u32, _u32, _vp]
        L.oracle_general_less_big.restype = C.c_int
        L.oracle_general_less_big.argtypes = [_vp, _u32, _u32, _u32, _u32]
        L.oracle_fnv1a_u32.restype = _u64
        L.oracle_fnv1a_u32.argtypes = [_vp, _u64]
        _oracle = L
    return _oracle


def have_ref():
    return os.path.exists(_REF_SO)


def ref():
    global _ref
    if _ref is None:
        L = C.CDLL(_REF_SO)
        for f in (L.ref_kiss1_suffix_array_dna, L.ref_kiss2_suffix_array_dna):
            f.restype = _dbl
            f.argtypes = [_vp, _u64, _u32, _u64, _vp]
        L.ref_kiss1_stages.restype = _u32
        L.ref_kiss1_stages.argtypes = [_vp, _u64, _u32, _u64, _vp, _vp, _vp, _vp, _vp, _vp]
        L.ref_induce_from_sorted_lms.restype = None
        L.ref_induce_from_sorted_lms.argtypes = [_vp, _u64, _u64, _vp, _vp]
        for name in ("ref_kiss1_suffix_array", "ref_kiss2_suffix_array"):
            if hasattr(L, name):
                getattr(L, name).restype = _dbl
                getattr(L, name).argtypes = [_vp, _u64, _u32, _u64, _vp]
        if hasattr(L, "ref_kiss2_sorted_lms"):
            L.ref_kiss2_sorted_lms.restype = _u32
            L.ref_kiss2_sorted_lms.argtypes = [_vp, _u64, _u32, _u64, _vp]
        if hasattr(L, "ref_fmindex_build_file"):
            L.ref_fmindex_build_file.restype = _u64
            L.ref_fmindex_build_file.argtypes = [_vp, _u64, C.c_char_p]
        if hasattr(L, "ref_read_sequence"):
            L.ref_read_sequence.restype = _u64
            L.ref_read_sequence.argtypes = [_vp, _u64, _vp]
        _ref = L
    return _ref


def _codes(S):
    S = np.ascontiguousarray(S, dtype=np.uint8)
    assert S.size == 0 or S.max() < 4
    return S


# ---------------------------------------------------------------- oracle (C restatement)
def oracle_kiss1(S, k=256, stages=False):
    S = _codes(S)
    n = S.size
    SA = np.zeros(n + 1, np.uint32)
    if not stages:
        oracle().oracle_kiss1_suffix_array_dna(_ptr(S), n, k & 0xFFFFFFFF, _ptr(SA), None, None, None, None, None)
        return SA
    lms = np.zeros(n // 2 + 2, np.uint32)
    srt = np.zeros(n // 2 + 2, np.uint32)
    placed = np.zeros(n + 1, np.uint32)
    cnt = np.zeros(4, np.uint32)
    lcnt = np.zeros(4, np.uint32)
    m = oracle().oracle_kiss1_suffix_array_dna(_ptr(S), n, k & 0xFFFFFFFF, _ptr(SA), _ptr(lms), _ptr(srt),
                                               _ptr(placed), _ptr(cnt), _ptr(lcnt))
    return dict(sa=SA, m=m, lms=lms[:m].copy(), sorted=srt[:m].copy(), placed=placed, count=cnt,
                lms_count=lcnt)


def oracle_get_lms(S):
    S = _codes(S)
    n = S.size
    lms = np.zeros(n // 2 + 2, np.uint32)
    cnt = np.zeros(4, np.uint32)
    lcnt = np.zeros(4, np.uint32)
    m = oracle().oracle_get_lms(_ptr(S), n, _ptr(lms), _ptr(cnt), _ptr(lcnt))
    return lms[:m].copy(), cnt, lcnt


def oracle_induce_from_sorted(S, sorted_lms):
    S = _codes(S)
    srt = np.ascontiguousarray(sorted_lms, dtype=np.uint32)
    SA = np.zeros(S.size + 1, np.uint32)
    oracle().oracle_induce_from_sorted_lms(_ptr(S), S.size, _ptr(srt), srt.size, _ptr(SA))
    return SA


def _bytes(S):
    if isinstance(S, str):
        S = S.encode("latin-1")
    if isinstance(S, (bytes, bytearray, memoryview)):
        return np.frombuffer(bytes(S), dtype=np.uint8)
    return np.ascontiguousarray(S, dtype=np.uint8)


def oracle_general(S, k=256):
    """general alphabet (oracle/general_oracle.c): LMS suffixes that share k symbols stay in text order"""
    S = _bytes(S)
    SA = np.zeros(S.size + 1, np.uint32)
    oracle().oracle_kiss1_suffix_array(_ptr(S) if S.size else None, S.size, k & 0xFFFFFFFF, _ptr(SA))
    return SA


def fnv1a(a):
    a = np.ascontiguousarray(a, dtype=np.uint32)
    return int(oracle().oracle_fnv1a_u32(_ptr(a), a.size))


# ---------------------------------------------------------------- the real reference
def ref_kiss1(S, k=256, threads=4):
    S = _codes(S)
    SA = np.zeros(S.size + 1, np.uint32)
    ref().ref_kiss1_suffix_array_dna(_ptr(S), S.size, k & 0xFFFFFFFF, threads, _ptr(SA))
    return SA


def ref_kiss2(S, k=256, threads=4):
    S = _codes(S)
    SA = np.zeros(S.size + 1, np.uint32)
    ref().ref_kiss2_suffix_array_dna(_ptr(S), S.size, k & 0xFFFFFFFF, threads, _ptr(SA))
    return SA


def ref_general(S, k=256, threads=4, algo=0):
    """the reference's general-alphabet kiss1_suffix_array / kiss2_suffix_array on bytes of any values"""
    S = _bytes(S)
    SA = np.zeros(S.size + 1, np.uint32)
    f = ref().ref_kiss1_suffix_array if algo == 0 else ref().ref_kiss2_suffix_array
    f(_ptr(S) if S.size else None, S.size, k & 0xFFFFFFFF, threads, _ptr(SA))
    return SA


def ref_kiss2_sorted_lms(S, k=256, threads=1):
    """kISS-2's LMS order (the reference's own encode_S .. place_back_lms, kiss2_core.hpp:728-798), sentinel first"""
    S = _codes(S)
    out = np.zeros(S.size // 2 + 2, np.uint32)
    m = ref().ref_kiss2_sorted_lms(_ptr(S), S.size, k & 0xFFFFFFFF, threads, _ptr(out))
    return out[:m].copy()


def ref_kiss1_stages(S, k=256, threads=4):
    S = _codes(S)
    n = S.size
    assert n > 0
    lms = np.zeros(n // 2 + 2, np.uint32)
    srt = np.zeros(n // 2 + 2, np.uint32)
    placed = np.zeros(n + 1, np.uint32)
    SA = np.zeros(n + 1, np.uint32)
    counters = np.zeros(5 * 256, np.uint32)
    sec = np.zeros(5, np.float64)
    m = ref().ref_kiss1_stages(_ptr(S), n, k & 0xFFFFFFFF, threads, _ptr(lms), _ptr(srt), _ptr(placed), _ptr(SA),
                               _ptr(counters), _ptr(sec))
    return dict(sa=SA, m=m, lms=lms[:m].copy(), sorted=srt[:m].copy(), placed=placed,
                count=counters[4 * 256:4 * 256 + 4].copy(), lms_count=counters[2 * 256:2 * 256 + 4].copy(),
                stage_sec=sec)


def naive_suffix_array(S):
    """O(n^2 log n) full suffix array with the sentinel first — for tiny n only."""
    S = bytes(_codes(S).tolist())
    n = len(S)
    return np.array([n] + sorted(range(n), key=lambda i: S[i:]), np.uint32)


def ref_fmindex_image(S):
    """the bytes of the .fmi file the reference's fmindex_build writes for the code text S (FMIndex<4, u32, KISS1Sorter>
    {.LOOKUP_LEN = 0}: build with k = 32, save; include/command/fmindex_build.hpp:27-34)"""
    import tempfile
    S = _codes(S)
    with tempfile.NamedTemporaryFile(suffix=".fmi") as f:
        ref().ref_fmindex_build_file(_ptr(S), S.size, f.name.encode())
        return open(f.name, "rb").read()


def ref_read_sequence(data):
    """the reference's own read_sequence (include/utils/io.hpp:6-18) + `% 4` on the bytes of a file"""
    data = bytes(data)
    buf = np.frombuffer(data, dtype=np.uint8) if data else np.zeros(0, np.uint8)
    out = np.zeros(max(len(data), 1), np.uint8)
    n = ref().ref_read_sequence(buf.ctypes.data if buf.size else None, len(data), out.ctypes.data)
    return out[:n].copy()


def sha_bytes(a):
    import hashlib
    return hashlib.sha256(np.ascontiguousarray(a, dtype=np.uint8).tobytes()).hexdigest()
