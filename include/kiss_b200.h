/*
 * kiss_b200.h — C ABI of the B200-native suffix_sort path (libkiss_b200.so).
 *
 * This is the drop-in boundary for jhhung/kISS's `suffix_sort` hot path.  Every entry point names the
 * reference interface it replaces (paths relative to the reference's include/biovoltron/algo/sort/
 * unless stated otherwise).  Plain pointers and sizes only; no CPU fallback exists behind any of
 * these: every call runs CUDA kernels on the selected device or returns an error code.
 *
 * Conventions
 *   - Text `S` is one code per byte, values 0..3 (A C G T), exactly what the reference's callers
 *     pass after `c % 4` (include/command/suffix_sort.hpp:33).  Only the low two bits are used.
 *   - Suffix array entries are uint32_t; `KISS_B200_EMPTY` (0xFFFFFFFF) is the reference's
 *     `EMPTY<uint32_t>` (constant.hpp:19-20), so n + 1 must be < 2^32 - 1.
 *   - k is the reference's `size_type k`; 0xFFFFFFFF means unbounded (`-k -1`,
 *     include/command/suffix_sort.hpp:35,59).
 *   - "host" pointers are ordinary host memory; "dev" pointers are device memory on the context's GPU.
 *   - All functions return 0 (KISS_B200_OK) or a negative error code; `kiss_b200_strerror` names it and
 *     `kiss_b200_last_error` gives the detailed message of the calling thread's last failure.
 */
#ifndef KISS_B200_H
#define KISS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KISS_B200_EMPTY 0xFFFFFFFFu

/* -s / --sorting-algorithm (reference: include/utils/options.hpp:67-81 SortingAlgorithm).
 * PREFIX_DOUBLING runs rank doubling over kISS-2's 16-base words (kiss2_core.hpp:728-798; csrc/pd.cu).  With
 * k = 0xFFFFFFFF the result is the unique suffix array; with a bounded k the loop stops at offset >= k like the
 * reference's (:764-766) and LMS suffixes whose first 2^ceil(log2 k) words are equal stay tied, in ascending text
 * position here — the reference leaves them in an order that depends on its thread count. */
#define KISS_B200_PARALLEL_SORTING 0
#define KISS_B200_PREFIX_DOUBLING 1

enum {
  KISS_B200_OK = 0,
  KISS_B200_ERR_ARG = -1,          /* null pointer, bad enum value */
  KISS_B200_ERR_TOO_LARGE = -2,    /* n + 1 >= 2^32 - 1 (EMPTY would collide, constant.hpp:19-20) */
  KISS_B200_ERR_CUDA = -3,         /* a CUDA runtime call or kernel failed */
  KISS_B200_ERR_NOMEM = -4,        /* device allocation failed (reference: std::bad_alloc, structs.hpp:32-38) */
  KISS_B200_ERR_UNSUPPORTED = -5,  /* (not returned any more: every k is accepted, like the reference) */
  KISS_B200_ERR_NO_DEVICE = -6     /* no CUDA device / the CUDA library cannot run: there is no CPU fallback */
};

typedef struct kiss_b200_ctx kiss_b200_ctx;

/* Per-stage device times (seconds, CUDA events on the context's stream), the same split the
 * reference logs with --verbose (kiss1_core.hpp:244-267; README.md:95-100). */
typedef struct kiss_b200_timings {
  double h2d;            /* host variant only: copy of S to the device */
  double pack;           /* 2-bit packing (reference: PackedDNAString ctor, structs.hpp:84-116) */
  double get_lms;        /* reference: get_lms, kiss_common.hpp:543-579 */
  double lms_sort;       /* reference: lms_suffix_direct_sort_dna, kiss1_core.hpp:23-145 */
  double put_lms;        /* reference: put_lms_suffix, kiss_common.hpp:445-481 */
  double induce;         /* reference: induced_sort, kiss_common.hpp:422-442 */
  double d2h;            /* host variant only: copy of SA back to the host; with a pinned destination the copies
                            of finished regions overlap the S pass and this is only the part left after it */
  double total_device;   /* pack .. induce */
  uint64_t n;            /* text length */
  uint64_t m;            /* LMS count including the sentinel (what get_lms returns, kiss_common.hpp:578) */
  uint64_t lms_sort_tied;   /* LMS suffixes still tied after the first 16-base radix key */
  uint64_t lms_sort_rounds; /* refinement rounds run for the large tied groups */
  uint64_t waves_L, waves_S; /* induction waves (maximal filled runs, kiss_common.hpp:389-393) */
  uint64_t kernel_launches; /* kernels launched by this call */
} kiss_b200_timings;

/* ---- context ----------------------------------------------------------------------------------
 * Owns the stream, the events and all device scratch; scratch is grown on demand and reused by
 * later calls (the reference allocates per call: kiss1_core.hpp:241-247).  One host thread per
 * context. */
int kiss_b200_create(kiss_b200_ctx** out, int device);
void kiss_b200_destroy(kiss_b200_ctx* ctx);
const char* kiss_b200_strerror(int code);
const char* kiss_b200_last_error(void);
int kiss_b200_abi_version(void);
/* CUDA stream (cudaStream_t) all work of this context is issued on */
void* kiss_b200_stream(kiss_b200_ctx* ctx);

/* ---- the path ---------------------------------------------------------------------------------
 * kiss_b200_suffix_sort_dna_u32: replaces
 *   KISS1Sorter<uint32_t>::get_suffix_array_dna(const kiss::vector<uint8_t>& S, uint32_t k, size_t threads)
 *     (kiss1_sorter.hpp:20-27) -> kiss::kiss1_suffix_array_dna (kiss1_core.hpp:229-268)   [algo 0]
 *   KISS2Sorter<uint32_t>::get_suffix_array_dna (kiss2_sorter.hpp:20-27)
 *     -> kiss::kiss2_suffix_array_dna (kiss2_core.hpp:888-896)                             [algo 1]
 * S_host: n codes; SA_host: n + 1 slots, filled with the suffix array (SA[0] = n); n == 0 gives {0}
 * (kiss1_core.hpp:237-238).  `t` may be NULL. */
int kiss_b200_suffix_sort_dna_u32(kiss_b200_ctx* ctx, const uint8_t* S_host, uint64_t n, uint32_t k,
                                  int algo, uint32_t* SA_host, kiss_b200_timings* t);

/* The same call over several GPUs of one node (the `n_gpus` argument; 8(e)): ctxs[0..n_gpus) are contexts
 * on the GPUs to use (one each; several contexts on one device also work and are what the 1-GPU tests use).  Only the LMS
 * sort shards — the text is uploaded to every GPU over its own PCIe link, rank r sorts the key range chosen by sampled
 * first-16-base splitters, LMS positions move by peer copies over NVLink, placement and induction run on ctxs[0] — and the
 * result is bit-identical to the single-GPU call.  One host thread per GPU inside the call; PREFIX_DOUBLING, k >= n - 1
 * and texts below 64 Ki bases run on ctxs[0] alone. */
int kiss_b200_suffix_sort_dna_u32_multi(kiss_b200_ctx** ctxs, int n_gpus, const uint8_t* S_host, uint64_t n, uint32_t k,
                                        int algo, uint32_t* SA_host, kiss_b200_timings* t);

/* Same, text and suffix array resident in device memory (no PCIe in the timed region).
 * S_dev must be 16-byte aligned. */
int kiss_b200_suffix_sort_dna_u32_dev(kiss_b200_ctx* ctx, const uint8_t* S_dev, uint64_t n, uint32_t k,
                                      int algo, uint32_t* SA_dev, kiss_b200_timings* t);

/* ASCII front end: bytes of sequence text (aA cC gG tT, anything else -> A) are coded on the device
 * exactly as Codec::to_int followed by `% 4` does (include/biovoltron/utility/istring.hpp:27-46,
 * include/command/suffix_sort.hpp:33), then sorted as above. */
int kiss_b200_suffix_sort_ascii_u32(kiss_b200_ctx* ctx, const char* seq_host, uint64_t n, uint32_t k,
                                    int algo, uint32_t* SA_host, kiss_b200_timings* t);

/* General alphabet: replaces
 *   KISS1Sorter<uint32_t>::get_suffix_array(const kiss::vector<uint8_t>& S, uint32_t k, size_t threads)
 *     (kiss1_sorter.hpp:38-44) -> kiss::kiss1_suffix_array (kiss1_core.hpp:270-311)        [algo 0]
 *   KISS2Sorter<uint32_t>::get_suffix_array (kiss2_sorter.hpp) -> kiss::kiss2_suffix_array  [algo 1]
 * S: n bytes of ANY values; SA: n + 1 slots (SA[0] = n; n == 0 gives {0}).  The text is reduced on the device to the
 * 2-bit alphabet — the byte values that occur are replaced by their ranks, written as b = 1..4 two-bit digits — and run
 * through the DNA path with k' = b * k; the entries at symbol boundaries are the result (general.cu).  b * n must stay
 * below 2^32 - 2 (KISS_B200_ERR_TOO_LARGE otherwise).
 * Order contract: every suffix is ordered by at least its first k symbols (the property tests/kiss.cpp:61-65 asserts);
 * k >= n - 1 gives the true suffix array.  Wherever the reference's result is unique — no two suffixes share their
 * first k symbols, or the true suffix array — the result is identical to it.  Among suffixes that share k symbols the
 * reference's own order is unspecified for n < 10^8 (an unstable std::sort on "first k symbols", kiss1_core.hpp:208-219);
 * here they are ordered deeper, as the DNA path orders the reduced text.
 * kiss_b200_general_digits reports the number of distinct byte values and b for a device-resident text. */
int kiss_b200_suffix_sort_u32(kiss_b200_ctx* ctx, const uint8_t* S_host, uint64_t n, uint32_t k, int algo,
                              uint32_t* SA_host, kiss_b200_timings* t);
int kiss_b200_suffix_sort_u32_dev(kiss_b200_ctx* ctx, const uint8_t* S_dev, uint64_t n, uint32_t k, int algo,
                                  uint32_t* SA_dev, kiss_b200_timings* t);
int kiss_b200_general_digits(kiss_b200_ctx* ctx, const uint8_t* S_dev, uint64_t n, uint32_t* sigma, uint32_t* digits);

/* ---- stage-level entries (used by the parity tests and the multi-GPU LMS sort) -------------------
 * Each one runs pack + the named stage(s) on the device and copies the stage's result to the host.
 *
 * kiss_b200_get_lms: reference get_lms (kiss_common.hpp:543-579).  lms_host receives the LMS positions
 * ascending followed by the sentinel n (capacity n/2 + 2); *m_out = their number including the sentinel;
 * count4 / lms_count4 receive the per-symbol totals the reference keeps in ThreadState::bucket
 * (kiss_common.hpp:519, 529-536).  Any output pointer may be NULL. */
int kiss_b200_get_lms(kiss_b200_ctx* ctx, const uint8_t* S_host, uint64_t n, uint32_t* lms_host,
                      uint64_t* m_out, uint64_t* count4, uint64_t* lms_count4);

/* kiss_b200_sorted_lms: get_lms + lms_suffix_direct_sort_dna (kiss1_core.hpp:23-145).  sorted_host
 * receives the m LMS suffixes in the reference's order, sorted_host[0] == n. */
int kiss_b200_sorted_lms(kiss_b200_ctx* ctx, const uint8_t* S_host, uint64_t n, uint32_t k, int algo,
                         uint32_t* sorted_host, uint64_t* m_out);

/* kiss_b200_put_lms: ... + put_lms_suffix (kiss_common.hpp:445-481): SA_host (n + 1) holds the sorted LMS
 * suffixes at their bucket ends, SA[0] = n, KISS_B200_EMPTY everywhere else. */
int kiss_b200_put_lms(kiss_b200_ctx* ctx, const uint8_t* S_host, uint64_t n, uint32_t k, int algo,
                      uint32_t* SA_host);

/* kiss_b200_induce_from_sorted_lms: put_lms_suffix + induced_sort (kiss_common.hpp:422-481) from a
 * caller-supplied order of the LMS suffixes (sorted_host[0] == n, m entries). */
int kiss_b200_induce_from_sorted_lms(kiss_b200_ctx* ctx, const uint8_t* S_host, uint64_t n,
                                     const uint32_t* sorted_host, uint64_t m, uint32_t* SA_host);

/* ---- device-resident stage API for the sharded LMS sort (one process per GPU) --------------------
 * kiss_b200_prepare_dev: pack + get_lms on this GPU for the whole text; keeps packed text and the
 * text-ordered LMS list in the context.  *m_out includes the sentinel.  S_dev must be 16-byte aligned. */
int kiss_b200_prepare_dev(kiss_b200_ctx* ctx, const uint8_t* S_dev, uint64_t n, uint32_t k, uint64_t* m_out);
/* Sort by the reference's LMS order an arbitrary subset of the long LMS suffixes given as device
 * positions in ascending text order (the shard a rank received); result in place. */
int kiss_b200_sort_lms_subset_dev(kiss_b200_ctx* ctx, uint32_t* pos_dev, uint64_t count, kiss_b200_timings* t);
/* First 16-base radix keys of the context's text-ordered long LMS suffixes (device pointers owned by the
 * context, valid until the next prepare): used to pick splitters and to bucket positions by destination. */
int kiss_b200_lms_keys_dev(kiss_b200_ctx* ctx, const uint32_t** keys_dev, const uint32_t** pos_dev,
                           uint64_t* n_long);
/* Stable partition of the context's text-ordered long LMS suffixes [lo, hi) by destination rank: destination of
 * a suffix = number of splitters <= its first-16-base key (n_split <= 63 ascending splitters, host array).
 * out_pos_dev receives the positions grouped by destination, ascending text order inside each group;
 * counts_host[0..n_split] the group sizes.  This is the send buffer of the NCCL all-to-all. */
int kiss_b200_partition_lms_dev(kiss_b200_ctx* ctx, const uint32_t* splitters_host, int n_split, uint64_t lo,
                                uint64_t hi, uint32_t* out_pos_dev, uint64_t* counts_host);
/* Finish on this GPU from a fully sorted list of the long LMS suffixes (device, n_long entries):
 * short-suffix insertion, placement and induction into SA_dev (n + 1). */
int kiss_b200_finish_from_sorted_dev(kiss_b200_ctx* ctx, const uint32_t* sorted_long_dev, uint64_t n_long,
                                     uint32_t* SA_dev, kiss_b200_timings* t);
/* Multi-GPU: the rank that sorted a key range also fetches the induction contexts of its suffixes (the 15 bases
 * before each position, one random 128-byte DRAM fetch per LMS suffix — the part of induced_L's first touch of an LMS
 * suffix, kiss_common.hpp:226-246, that does shard), and rank 0 finishes from (position, context) pairs.
 * lms_contexts: ctx_out_dev[i] for the suffixes pos_dev[0..count) of the prepared text; the call returns when the
 * kernel has finished.
 * finish_from_sorted_ctx: as finish_from_sorted_dev; ctx_long_dev (nullable) is aligned like sorted_long_dev mod 16. */
int kiss_b200_lms_contexts_dev(kiss_b200_ctx* ctx, const uint32_t* pos_dev, uint64_t count, uint32_t* ctx_out_dev);
int kiss_b200_finish_from_sorted_ctx_dev(kiss_b200_ctx* ctx, const uint32_t* sorted_long_dev, const uint32_t* ctx_long_dev,
                                         uint64_t n_long, uint32_t* SA_dev, kiss_b200_timings* t);

/* ---- ingest: the step before the path ------------------------------------
 * The bytes of a FASTA or plain-text file -> base codes, on the device, with the semantics of the reference's
 * read_sequence (include/utils/io.hpp:6-18; FASTA extraction include/biovoltron/file_io/fasta.hpp:116-151; Codec::to_int
 * + `% 4`, istring.hpp:27-46, include/command/suffix_sort.hpp:33): FASTA mode iff the first byte is '>', header lines
 * (first byte '>') dropped whole, '\n' dropped, everything else kept — a/A c/C g/G t/T -> 0..3, any other byte (N,
 * IUPAC codes, the '\r' of a CRLF file) -> 0 — and all records concatenated.  The reference does this with serial
 * getline + push_back; here it is a stream compaction in three kernels.
 * suffix_sort_file: file bytes on the host in, suffix array out; *n_out = text length; SA_host holds SA_capacity
 *   entries (file size + 1 always suffices); KISS_B200_ERR_ARG when the text turns out longer than SA_capacity - 1.
 * ingest_dev: file_dev[0..nbytes) -> codes_dev (room for nbytes bytes), *n_out = text length; fasta = first byte is '>'. */
int kiss_b200_suffix_sort_file_u32(kiss_b200_ctx* ctx, const uint8_t* file_host, uint64_t nbytes, uint32_t k, int algo,
                                   uint32_t* SA_host, uint64_t SA_capacity, uint64_t* n_out, kiss_b200_timings* t);
int kiss_b200_ingest_dev(kiss_b200_ctx* ctx, const uint8_t* file_dev, uint64_t nbytes, int fasta, uint8_t* codes_dev,
                         uint64_t* n_out);

/* ---- FM-index construction: the step after the path -----------------
 * The reference's `fmindex_build` command (include/command/fmindex_build.hpp:27-34) builds
 * FMIndex<SA_INTV = 4, uint32_t, KISS1Sorter<uint32_t>>{.LOOKUP_LEN = 0}: suffix array with k = 32 (fm_index.hpp:385-386),
 * then build_occ (fm_index.hpp:278-307), build_bwt (:311-328), build_sa (:332-372), build_lookup (:237-269; {0, n + 1}
 * for LOOKUP_LEN = 0), and writes it with FMIndex::save (:591-617) through Serializer::save
 * (include/biovoltron/utility/archive/serializer.hpp:95-112).  Here the suffix array stays in device memory: the BWT,
 * both occ tables, the sampled-entry bit vector with its prefix sums and the sampled suffix array are built by kernels
 * (csrc/fmindex.cu) and only they cross PCIe (1.75 bytes per suffix instead of 4), laid out in `image_host` as exactly
 * the bytes of the reference's .fmi file — its `fmindex_query` loads the result.  (The reference leaves the bits past the
 * last element in the last byte of bwt_ and the last word of b_ uninitialised; they are zero here.)
 * sa_intv = the reference's SA_INTV template argument (4 in its command; 1 stores the whole array and no bit vector).
 * fmindex_image_bytes: size of the image for a text of n bases.
 * fmindex_build_dna: S_host = n codes 0..3; k / algo as kiss_b200_suffix_sort_dna_u32 (the reference uses 32 / kISS-1).
 * fmindex_build_file: the bytes of a FASTA / text file in (read_sequence semantics, see the ingest section); *n_out =
 *   text length; capacity kiss_b200_fmindex_image_bytes(nbytes, sa_intv) always suffices.
 * fmindex_from_sa_dev: from a suffix array already in device memory (S_dev: codes, 16-byte aligned; SA_dev: n + 1).
 * *image_bytes = bytes written; KISS_B200_ERR_ARG when capacity is too small.  t, fm_seconds (device seconds of the
 * FM-index kernels) may be NULL. */
uint64_t kiss_b200_fmindex_image_bytes(uint64_t n, uint32_t sa_intv);
int kiss_b200_fmindex_build_dna_u32(kiss_b200_ctx* ctx, const uint8_t* S_host, uint64_t n, uint32_t k, int algo,
                                    uint32_t sa_intv, uint8_t* image_host, uint64_t capacity, uint64_t* image_bytes,
                                    kiss_b200_timings* t, double* fm_seconds);
int kiss_b200_fmindex_build_file_u32(kiss_b200_ctx* ctx, const uint8_t* file_host, uint64_t nbytes, uint32_t k, int algo,
                                     uint32_t sa_intv, uint8_t* image_host, uint64_t capacity, uint64_t* image_bytes,
                                     uint64_t* n_out, kiss_b200_timings* t, double* fm_seconds);
int kiss_b200_fmindex_from_sa_dev(kiss_b200_ctx* ctx, const uint8_t* S_dev, uint64_t n, const uint32_t* SA_dev,
                                  uint32_t sa_intv, uint8_t* image_host, uint64_t capacity, uint64_t* image_bytes,
                                  double* fm_seconds);

/* ---- allocation-free ingest ---------------------------------------------------------------------
 * kiss_b200_ingest_noalloc_dev has NO COUNTERPART IN THE REFERENCE: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same arguments, kernels and result as kiss_b200_ingest_dev, but the per-chunk scratch lives in a file-scope
 * __device__ array of the library (about 0.66 MB on every GPU that loads it, held until the process exits) instead of
 * the context's arena: the call makes no cudaMalloc / cudaFree / allocating library call.  Limit: nbytes <=
 * KISS_B200_NOALLOC_MAX_BYTES, else KISS_B200_ERR_TOO_LARGE (kiss_b200_last_error says so).  Every call on a GPU shares
 * the one array — it keeps the last call's contents; nothing in it has to be cleared by the caller — so calls that
 * overlap on one GPU (two contexts, two host threads) race: serialise them. */
#define KISS_B200_NOALLOC_MAX_BYTES (256ull << 20)
int kiss_b200_ingest_noalloc_dev(kiss_b200_ctx* ctx, const uint8_t* file_dev, uint64_t nbytes, int fasta,
                                 uint8_t* codes_dev, uint64_t* n_out);

/* ---- per-kernel timing ---------------------------------------------------------------------------
 * When enabled, every launch of the main kernel classes is bracketed by CUDA events on the context's stream and
 * the durations are accumulated per class (bench.py's roofline numbers).  enable() also clears the totals.
 * profile_read: idx = 0, 1, ... until it returns KISS_B200_ERR_ARG. */
int kiss_b200_profile_enable(kiss_b200_ctx* ctx, int on);
int kiss_b200_profile_read(kiss_b200_ctx* ctx, int idx, const char** name, double* ms, uint64_t* launches);
/* algorithmic bytes of the launches of class idx, where the launcher knows them (radix passes: keys and positions
 * read + written; tie-group classes: 16 B per LMS suffix in the groups refined); else 0 */
int kiss_b200_profile_alg_bytes(kiss_b200_ctx* ctx, int idx, uint64_t* bytes);

/* kernels launched so far by the calling host thread through this library (a counter, not a CUDA query): callers that
 * drive the device-stage entries themselves (kiss_b200/multi_gpu.py) report the difference as their launch count */
uint64_t kiss_b200_launch_count(void);

/* ---- result digest (parity at sizes where the array itself cannot be kept: SURVEY.md 8(c), tests/golden) -----------
 * *out = sum over i < count of splitmix64((i << 32) | SA_dev[i]) mod 2^64: an order-sensitive checksum of a suffix array
 * that is a plain sum, so one streaming pass computes it on the device.  tests/golden/full_size_golden.json holds the
 * value for the REFERENCE's own suffix array (kiss1_suffix_array_dna, kiss1_core.hpp:229-268) of the BASELINE.json
 * configurations; bench.py compares after every run, outside the timed region. */
int kiss_b200_sa_checksum_dev(kiss_b200_ctx* ctx, const uint32_t* SA_dev, uint64_t count, uint64_t* out);

/* The property the reference's own tests assert (tests/kiss.cpp:25-33: REQUIRE(substr(sa[i-1], k) <= substr(sa[i], k))),
 * checked on the device over the whole array, plus SA[0] == n and "SA is a permutation of 0..n".  No counterpart in the
 * reference's library (its check is that test loop).  S_dev: n codes, 16-byte aligned; SA_dev: n + 1 entries.
 * counts[0] = adjacent pairs in the wrong order, [1] = values that occur twice, [2] = values out of range (or SA[0] != n),
 * [3] = pairs whose comparison was cut after 4096 equal 32-base words (only with k beyond 131 072 on repetitive text).
 * A correct k-ordered suffix array gives {0, 0, 0, 0}. */
int kiss_b200_verify_sa_dev(kiss_b200_ctx* ctx, const uint8_t* S_dev, uint64_t n, const uint32_t* SA_dev, uint32_t k,
                            uint64_t counts[4]);
/* The same for a byte text and the array of the general-alphabet entry (symbols compared as unsigned bytes; [3] counts
 * pairs cut after 131 072 equal symbols). */
int kiss_b200_verify_sa_bytes_dev(kiss_b200_ctx* ctx, const uint8_t* S_dev, uint64_t n, const uint32_t* SA_dev, uint32_t k,
                                  uint64_t counts[4]);

/* ---- synthetic inputs ---------------------------------------------------------
 * Fill S_dev[0..n) with the counter-hash uniform stream `seed` (identical to tests/textgen.py uniform). */
int kiss_b200_synth_uniform_dev(kiss_b200_ctx* ctx, uint8_t* S_dev, uint64_t n, uint64_t seed);

#ifdef __cplusplus
}
#endif
#endif /* KISS_B200_H */
