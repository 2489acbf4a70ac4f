"""ctypes binding of libkiss_b200.so (include/kiss_b200.h).  Loading never falls back to anything:
if the library is missing or no CUDA device can be opened, the caller gets an exception."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libkiss_b200.so")

PARALLEL_SORTING = 0
PREFIX_DOUBLING = 1
EMPTY = 0xFFFFFFFF
UNBOUNDED_K = 0xFFFFFFFF
NOALLOC_MAX_BYTES = 256 << 20  # KISS_B200_NOALLOC_MAX_BYTES


class Timings(C.Structure):
    _fields_ = [(n, C.c_double) for n in
                ("h2d", "pack", "get_lms", "lms_sort", "put_lms", "induce", "d2h", "total_device")] + \
               [(n, C.c_uint64) for n in
                ("n", "m", "lms_sort_tied", "lms_sort_rounds", "waves_L", "waves_S", "kernel_launches")]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class KissB200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"kiss_b200 error {code}: {msg}")
        self.code = code


_lib = None
_vp, _u32, _u64, _int = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int

# every symbol include/kiss_b200.h declares: (restype, argtypes)
SIGNATURES = {
    "kiss_b200_create": (_int, [C.POINTER(_vp), _int]),
    "kiss_b200_destroy": (None, [_vp]),
    "kiss_b200_strerror": (C.c_char_p, [_int]),
    "kiss_b200_last_error": (C.c_char_p, []),
    "kiss_b200_abi_version": (_int, []),
    "kiss_b200_stream": (_vp, [_vp]),
    "kiss_b200_suffix_sort_dna_u32": (_int, [_vp, _vp, _u64, _u32, _int, _vp, C.POINTER(Timings)]),
    "kiss_b200_suffix_sort_dna_u32_multi": (_int, [C.POINTER(_vp), _int, _vp, _u64, _u32, _int, _vp, C.POINTER(Timings)]),
    "kiss_b200_suffix_sort_dna_u32_dev": (_int, [_vp, _vp, _u64, _u32, _int, _vp, C.POINTER(Timings)]),
    "kiss_b200_suffix_sort_ascii_u32": (_int, [_vp, _vp, _u64, _u32, _int, _vp, C.POINTER(Timings)]),
    "kiss_b200_suffix_sort_u32": (_int, [_vp, _vp, _u64, _u32, _int, _vp, C.POINTER(Timings)]),
    "kiss_b200_suffix_sort_u32_dev": (_int, [_vp, _vp, _u64, _u32, _int, _vp, C.POINTER(Timings)]),
    "kiss_b200_general_digits": (_int, [_vp, _vp, _u64, C.POINTER(_u32), C.POINTER(_u32)]),
    "kiss_b200_verify_sa_dev": (_int, [_vp, _vp, _u64, _vp, _u32, _vp]),
    "kiss_b200_verify_sa_bytes_dev": (_int, [_vp, _vp, _u64, _vp, _u32, _vp]),
    "kiss_b200_get_lms": (_int, [_vp, _vp, _u64, _vp, C.POINTER(_u64), _vp, _vp]),
    "kiss_b200_sorted_lms": (_int, [_vp, _vp, _u64, _u32, _int, _vp, C.POINTER(_u64)]),
    "kiss_b200_put_lms": (_int, [_vp, _vp, _u64, _u32, _int, _vp]),
    "kiss_b200_induce_from_sorted_lms": (_int, [_vp, _vp, _u64, _vp, _u64, _vp]),
    "kiss_b200_prepare_dev": (_int, [_vp, _vp, _u64, _u32, C.POINTER(_u64)]),
    "kiss_b200_sort_lms_subset_dev": (_int, [_vp, _vp, _u64, C.POINTER(Timings)]),
    "kiss_b200_partition_lms_dev": (_int, [_vp, _vp, _int, _u64, _u64, _vp, _vp]),
    "kiss_b200_lms_keys_dev": (_int, [_vp, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_u64)]),
    "kiss_b200_finish_from_sorted_dev": (_int, [_vp, _vp, _u64, _vp, C.POINTER(Timings)]),
    "kiss_b200_finish_from_sorted_ctx_dev": (_int, [_vp, _vp, _vp, _u64, _vp, C.POINTER(Timings)]),
    "kiss_b200_lms_contexts_dev": (_int, [_vp, _vp, _u64, _vp]),
    "kiss_b200_suffix_sort_file_u32": (_int, [_vp, _vp, _u64, _u32, _int, _vp, _u64, C.POINTER(_u64), C.POINTER(Timings)]),
    "kiss_b200_ingest_dev": (_int, [_vp, _vp, _u64, _int, _vp, C.POINTER(_u64)]),
    "kiss_b200_fmindex_image_bytes": (_u64, [_u64, _u32]),
    "kiss_b200_fmindex_build_dna_u32": (_int, [_vp, _vp, _u64, _u32, _int, _u32, _vp, _u64, C.POINTER(_u64),
                                               C.POINTER(Timings), C.POINTER(C.c_double)]),
    "kiss_b200_fmindex_build_file_u32": (_int, [_vp, _vp, _u64, _u32, _int, _u32, _vp, _u64, C.POINTER(_u64),
                                                C.POINTER(_u64), C.POINTER(Timings), C.POINTER(C.c_double)]),
    "kiss_b200_fmindex_from_sa_dev": (_int, [_vp, _vp, _u64, _vp, _u32, _vp, _u64, C.POINTER(_u64), C.POINTER(C.c_double)]),
    "kiss_b200_ingest_noalloc_dev": (_int, [_vp, _vp, _u64, _int, _vp, C.POINTER(_u64)]),
    "kiss_b200_profile_enable": (_int, [_vp, _int]),
    "kiss_b200_profile_read": (_int, [_vp, _int, C.POINTER(C.c_char_p), C.POINTER(C.c_double), C.POINTER(_u64)]),
    "kiss_b200_profile_alg_bytes": (_int, [_vp, _int, C.POINTER(_u64)]),
    "kiss_b200_launch_count": (_u64, []),
    "kiss_b200_sa_checksum_dev": (_int, [_vp, _vp, _u64, C.POINTER(_u64)]),
    "kiss_b200_synth_uniform_dev": (_int, [_vp, _vp, _u64, _u64]),
}


def lib():
    """the loaded library; raises if it has not been built (python __graft_entry__.py build)"""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise KissB200Error(-6, f"{LIB_PATH} is not built; run `make -C kiss_b200/csrc` "
                                    "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            f = getattr(L, name)
            f.restype = res
            f.argtypes = args
        _lib = L
    return _lib


def check(code):
    if code != 0:
        L = lib()
        detail = L.kiss_b200_last_error().decode(errors="replace")
        raise KissB200Error(code, f"{L.kiss_b200_strerror(code).decode()}: {detail}")
