"""Host-side mirror of the reference's sorter facades over the C ABI.

Reference interface (include/biovoltron/algo/sort/kiss1_sorter.hpp:7-50):

    template<typename size_type = std::uint32_t> struct KISS1Sorter {
      static auto get_suffix_array_dna(const kiss::vector<uint8_t>& S, size_type k = 256u,
                                       size_t num_threads = hardware_concurrency()) -> SA_t;
      static auto prepare_aligned_ref(istring_view ref) -> kiss::vector<uint8_t>;
    };

Same names and argument meaning here; `num_threads` is accepted and ignored (the GPU path has no host
threads, and the reference's result does not depend on it).  Errors the reference raises as
exceptions (std::bad_alloc, structs.hpp:32-38) surface as KissB200Error.
"""
import ctypes as C

import numpy as np

from . import _cabi
from ._cabi import PARALLEL_SORTING, PREFIX_DOUBLING, Timings, check, lib


def _codes(S):
    """what every reference caller does before the sorter: codes 0..3, contiguous bytes
    (include/command/suffix_sort.hpp:33, kiss1_sorter.hpp:46-49)"""
    if isinstance(S, (bytes, bytearray, memoryview)):
        S = np.frombuffer(S, dtype=np.uint8)
    S = np.ascontiguousarray(S, dtype=np.uint8)
    return S


class Context:
    """one GPU context (stream + cached device scratch); not thread-safe, like the reference's call"""

    def __init__(self, device=0):
        self._h = C.c_void_p()
        check(lib().kiss_b200_create(C.byref(self._h), int(device)))
        self.device = device
        self.last_timings = None

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib().kiss_b200_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def handle(self):
        return self._h

    # ---- the path
    def suffix_array_dna(self, S, k=256, algo=PARALLEL_SORTING, out=None):
        S = _codes(S)
        n = S.size
        SA = np.empty(n + 1, np.uint32) if out is None else out
        t = Timings()
        check(lib().kiss_b200_suffix_sort_dna_u32(self._h, S.ctypes.data, n, k & 0xFFFFFFFF, algo, SA.ctypes.data,
                                                  C.byref(t)))
        self.last_timings = t.as_dict()
        return SA

    def suffix_array_ascii(self, seq, k=256, algo=PARALLEL_SORTING):
        b = np.frombuffer(seq if isinstance(seq, (bytes, bytearray)) else seq.encode(), dtype=np.uint8)
        b = np.ascontiguousarray(b)
        SA = np.empty(b.size + 1, np.uint32)
        t = Timings()
        check(lib().kiss_b200_suffix_sort_ascii_u32(self._h, b.ctypes.data, b.size, k & 0xFFFFFFFF, algo,
                                                    SA.ctypes.data, C.byref(t)))
        self.last_timings = t.as_dict()
        return SA

    def suffix_array(self, S, k=256, algo=PARALLEL_SORTING):
        """general alphabet: KISS1Sorter / KISS2Sorter ::get_suffix_array (kiss1_sorter.hpp:28-44).  `S`: bytes, str or
        a uint8 array of any values.  See include/kiss_b200.h for the order among suffixes that share k symbols."""
        if isinstance(S, str):
            S = S.encode("latin-1")
        b = np.frombuffer(S, dtype=np.uint8) if isinstance(S, (bytes, bytearray, memoryview)) else np.asarray(S)
        if b.dtype != np.uint8:
            raise TypeError("general-alphabet text must be bytes or a uint8 array")
        b = np.ascontiguousarray(b)
        SA = np.empty(b.size + 1, np.uint32)
        t = Timings()
        check(lib().kiss_b200_suffix_sort_u32(self._h, b.ctypes.data, b.size, k & 0xFFFFFFFF, algo, SA.ctypes.data,
                                              C.byref(t)))
        self.last_timings = t.as_dict()
        return SA

    def suffix_array_dev(self, S_ptr, n, SA_ptr, k=256, algo=PARALLEL_SORTING):
        """general alphabet, text and result resident in device memory"""
        t = Timings()
        check(lib().kiss_b200_suffix_sort_u32_dev(self._h, S_ptr, n, k & 0xFFFFFFFF, algo, SA_ptr, C.byref(t)))
        self.last_timings = t.as_dict()
        return self.last_timings

    def verify_sa_dev(self, S_ptr, n, SA_ptr, k=256, byte_text=False):
        """tests/kiss.cpp's REQUIRE loop on the device: dict(order_violations, duplicates, out_of_range, cut) — all zero for
        a correct k-ordered suffix array.  `byte_text`: S is a general-alphabet text (any byte values), not codes 0..3"""
        out = np.zeros(4, np.uint64)
        f = lib().kiss_b200_verify_sa_bytes_dev if byte_text else lib().kiss_b200_verify_sa_dev
        check(f(self._h, S_ptr, n, SA_ptr, k & 0xFFFFFFFF, out.ctypes.data))
        return dict(order_violations=int(out[0]), duplicates=int(out[1]), out_of_range=int(out[2]), cut=int(out[3]))

    def general_digits(self, S_ptr, n):
        """(distinct byte values, two-bit digits per symbol) of a device-resident text"""
        s, d = C.c_uint32(), C.c_uint32()
        check(lib().kiss_b200_general_digits(self._h, S_ptr, n, C.byref(s), C.byref(d)))
        return s.value, d.value

    def suffix_array_file(self, data, k=256, algo=PARALLEL_SORTING):
        """`data` = the bytes of a FASTA / plain-text file (read_sequence semantics, include/utils/io.hpp:6-18):
        returns (suffix array, text length)"""
        b = np.frombuffer(data, dtype=np.uint8) if isinstance(data, (bytes, bytearray, memoryview)) else data
        b = np.ascontiguousarray(b, dtype=np.uint8)
        SA = np.empty(b.size + 1, np.uint32)
        n, t = C.c_uint64(), Timings()
        check(lib().kiss_b200_suffix_sort_file_u32(self._h, b.ctypes.data, b.size, k & 0xFFFFFFFF, algo, SA.ctypes.data,
                                                   SA.size, C.byref(n), C.byref(t)))
        self.last_timings = t.as_dict()
        return SA[:n.value + 1].copy(), n.value

    def ingest_dev(self, file_ptr, nbytes, fasta, codes_ptr):
        n = C.c_uint64()
        check(lib().kiss_b200_ingest_dev(self._h, file_ptr, nbytes, 1 if fasta else 0, codes_ptr, C.byref(n)))
        return n.value

    # ---- FM-index (the step after the path): the reference's .fmi image, the suffix array never leaves the device
    def fmindex_build(self, S, k=32, algo=PARALLEL_SORTING, sa_intv=4, out=None):
        """FMIndex<sa_intv, uint32_t, Sorter>{.LOOKUP_LEN = 0}.build(S) + save (fm_index.hpp:380-411, 591-617): the
        bytes of the .fmi file (numpy uint8; `out`: a preallocated, e.g. pinned, buffer)"""
        S = _codes(S)
        need = lib().kiss_b200_fmindex_image_bytes(S.size, sa_intv)
        img = np.empty(need, np.uint8) if out is None else out
        nb, t, fm = C.c_uint64(), Timings(), C.c_double()
        check(lib().kiss_b200_fmindex_build_dna_u32(self._h, S.ctypes.data, S.size, k & 0xFFFFFFFF, algo, sa_intv,
                                                    img.ctypes.data, img.size, C.byref(nb), C.byref(t), C.byref(fm)))
        self.last_timings = dict(t.as_dict(), fmindex=fm.value)
        return img[:nb.value]

    def fmindex_build_file(self, data, k=32, algo=PARALLEL_SORTING, sa_intv=4):
        """`data` = the bytes of a FASTA / text file: (.fmi image, text length)"""
        b = np.frombuffer(data, dtype=np.uint8) if isinstance(data, (bytes, bytearray, memoryview)) else data
        b = np.ascontiguousarray(b, dtype=np.uint8)
        img = np.empty(lib().kiss_b200_fmindex_image_bytes(b.size, sa_intv), np.uint8)
        nb, n, t, fm = C.c_uint64(), C.c_uint64(), Timings(), C.c_double()
        check(lib().kiss_b200_fmindex_build_file_u32(self._h, b.ctypes.data, b.size, k & 0xFFFFFFFF, algo, sa_intv,
                                                     img.ctypes.data, img.size, C.byref(nb), C.byref(n), C.byref(t),
                                                     C.byref(fm)))
        self.last_timings = dict(t.as_dict(), fmindex=fm.value)
        return img[:nb.value].copy(), n.value

    def fmindex_from_sa_dev(self, S_ptr, n, SA_ptr, sa_intv=4, out=None):
        need = lib().kiss_b200_fmindex_image_bytes(n, sa_intv)
        img = np.empty(need, np.uint8) if out is None else out
        nb, fm = C.c_uint64(), C.c_double()
        check(lib().kiss_b200_fmindex_from_sa_dev(self._h, S_ptr, n, SA_ptr, sa_intv, img.ctypes.data, img.size,
                                                  C.byref(nb), C.byref(fm)))
        self.last_fm_seconds = fm.value
        return img[:nb.value]

    def ingest_noalloc_dev(self, file_ptr, nbytes, fasta, codes_ptr):
        """ingest_dev with its scratch in a fixed __device__ array: no allocation, at most NOALLOC_MAX_BYTES"""
        n = C.c_uint64()
        check(lib().kiss_b200_ingest_noalloc_dev(self._h, file_ptr, nbytes, 1 if fasta else 0, codes_ptr, C.byref(n)))
        return n.value

    def suffix_array_dna_dev(self, S_ptr, n, SA_ptr, k=256, algo=PARALLEL_SORTING):
        """device-resident variant: S_ptr / SA_ptr are device addresses (e.g. torch tensor.data_ptr())"""
        t = Timings()
        check(lib().kiss_b200_suffix_sort_dna_u32_dev(self._h, S_ptr, n, k & 0xFFFFFFFF, algo, SA_ptr, C.byref(t)))
        self.last_timings = t.as_dict()
        return self.last_timings

    # ---- stage-level entries (parity tests)
    def get_lms(self, S):
        S = _codes(S)
        n = S.size
        lms = np.empty(n // 2 + 2, np.uint32)
        m = C.c_uint64()
        cnt = np.zeros(4, np.uint64)
        lcnt = np.zeros(4, np.uint64)
        check(lib().kiss_b200_get_lms(self._h, S.ctypes.data, n, lms.ctypes.data, C.byref(m), cnt.ctypes.data,
                                      lcnt.ctypes.data))
        return lms[:m.value].copy(), cnt, lcnt

    def sorted_lms(self, S, k=256, algo=PARALLEL_SORTING):
        S = _codes(S)
        n = S.size
        out = np.empty(n // 2 + 2, np.uint32)
        m = C.c_uint64()
        check(lib().kiss_b200_sorted_lms(self._h, S.ctypes.data, n, k & 0xFFFFFFFF, algo, out.ctypes.data, C.byref(m)))
        return out[:m.value].copy()

    def put_lms(self, S, k=256, algo=PARALLEL_SORTING):
        S = _codes(S)
        SA = np.empty(S.size + 1, np.uint32)
        check(lib().kiss_b200_put_lms(self._h, S.ctypes.data, S.size, k & 0xFFFFFFFF, algo, SA.ctypes.data))
        return SA

    def induce_from_sorted_lms(self, S, sorted_lms):
        S = _codes(S)
        srt = np.ascontiguousarray(sorted_lms, dtype=np.uint32)
        SA = np.empty(S.size + 1, np.uint32)
        check(lib().kiss_b200_induce_from_sorted_lms(self._h, S.ctypes.data, S.size, srt.ctypes.data, srt.size,
                                                     SA.ctypes.data))
        return SA

    # ---- device-resident stage API (sharded LMS sort)
    def prepare_dev(self, S_ptr, n, k=256):
        m = C.c_uint64()
        check(lib().kiss_b200_prepare_dev(self._h, S_ptr, n, k & 0xFFFFFFFF, C.byref(m)))
        return m.value

    def lms_keys_dev(self):
        kp, pp, nl = C.c_void_p(), C.c_void_p(), C.c_uint64()
        check(lib().kiss_b200_lms_keys_dev(self._h, C.byref(kp), C.byref(pp), C.byref(nl)))
        return kp.value, pp.value, nl.value

    def partition_lms_dev(self, splitters, lo, hi, out_ptr):
        spl = np.ascontiguousarray(splitters, dtype=np.uint32)
        counts = np.zeros(spl.size + 1, np.uint64)
        check(lib().kiss_b200_partition_lms_dev(self._h, spl.ctypes.data, spl.size, lo, hi, out_ptr, counts.ctypes.data))
        return counts

    def sort_lms_subset_dev(self, pos_ptr, count):
        t = Timings()
        check(lib().kiss_b200_sort_lms_subset_dev(self._h, pos_ptr, count, C.byref(t)))
        return t.as_dict()

    def finish_from_sorted_dev(self, sorted_ptr, n_long, SA_ptr, ctx_ptr=None):
        t = Timings()
        if ctx_ptr:
            check(lib().kiss_b200_finish_from_sorted_ctx_dev(self._h, sorted_ptr, ctx_ptr, n_long, SA_ptr, C.byref(t)))
        else:
            check(lib().kiss_b200_finish_from_sorted_dev(self._h, sorted_ptr, n_long, SA_ptr, C.byref(t)))
        return t.as_dict()

    def lms_contexts_dev(self, pos_ptr, count, out_ptr):
        check(lib().kiss_b200_lms_contexts_dev(self._h, pos_ptr, count, out_ptr))

    def profile_enable(self, on=True):
        check(lib().kiss_b200_profile_enable(self._h, 1 if on else 0))

    def profile(self):
        """{kernel class: (total ms, launches)} accumulated since profile_enable()"""
        out, i = {}, 0
        name, ms, cnt = C.c_char_p(), C.c_double(), C.c_uint64()
        while lib().kiss_b200_profile_read(self._h, i, C.byref(name), C.byref(ms), C.byref(cnt)) == 0:
            if cnt.value:
                out[name.value.decode()] = (ms.value, cnt.value)
            i += 1
        return out

    def profile_alg_bytes(self):
        """{kernel class: algorithmic bytes of its launches} for the classes whose launcher knows them"""
        out, i = {}, 0
        name, ms, cnt, units = C.c_char_p(), C.c_double(), C.c_uint64(), C.c_uint64()
        while lib().kiss_b200_profile_read(self._h, i, C.byref(name), C.byref(ms), C.byref(cnt)) == 0:
            check(lib().kiss_b200_profile_alg_bytes(self._h, i, C.byref(units)))
            if units.value:
                out[name.value.decode()] = units.value
            i += 1
        return out

    def sa_checksum_dev(self, SA_ptr, count):
        """order-sensitive checksum of a device suffix array (sum of splitmix64((i << 32) | SA[i]) mod 2^64)"""
        out = C.c_uint64()
        check(lib().kiss_b200_sa_checksum_dev(self._h, SA_ptr, count, C.byref(out)))
        return out.value

    def synth_uniform_dev(self, S_ptr, n, seed):
        check(lib().kiss_b200_synth_uniform_dev(self._h, S_ptr, n, seed))


def suffix_array_dna_multi(contexts, S, k=256, algo=PARALLEL_SORTING):
    """kiss_b200_suffix_sort_dna_u32_multi: the sharded LMS sort over the GPUs of `contexts` (a list of Context), host
    buffers in and out; returns (suffix array, timings dict)"""
    S = _codes(S)
    SA = np.empty(S.size + 1, np.uint32)
    arr = (C.c_void_p * len(contexts))(*[c.handle.value for c in contexts])
    t = Timings()
    check(lib().kiss_b200_suffix_sort_dna_u32_multi(arr, len(contexts), S.ctypes.data, S.size, k & 0xFFFFFFFF, algo,
                                                    SA.ctypes.data, C.byref(t)))
    return SA, t.as_dict()


_default = None


def default_context():
    global _default
    if _default is None:
        _default = Context(0)
    return _default


class KISS1Sorter:
    """kiss1_sorter.hpp:7-50 (PARALLEL_SORTING)"""
    algo = PARALLEL_SORTING

    @classmethod
    def get_suffix_array_dna(cls, S, k=256, num_threads=None):
        return default_context().suffix_array_dna(S, k, cls.algo)

    @classmethod
    def get_suffix_array(cls, S, k=256, num_threads=None):
        """general alphabet (kiss1_sorter.hpp:28-44)"""
        return default_context().suffix_array(S, k, cls.algo)

    @staticmethod
    def prepare_aligned_ref(ref):
        return _codes(ref).copy()


class KISS2Sorter(KISS1Sorter):
    """kiss2_sorter.hpp:7-50 (PREFIX_DOUBLING)"""
    algo = PREFIX_DOUBLING
