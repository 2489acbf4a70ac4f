// kernels.h — host-side launchers of the sm_100a kernels (internal; the public boundary is
// include/kiss_b200.h).  Every launcher enqueues on `st` and never synchronises unless it says so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <stdexcept>
#include <string>
#include <vector>

typedef unsigned long long u64;
typedef unsigned int u32;
typedef unsigned char u8;

struct kiss_b200_ctx;

namespace kiss {

// api.cu, for the entry points defined in other files
void api_set_last_error(const std::string& msg);
int api_ctx_device(const kiss_b200_ctx* c);

struct CudaError : std::runtime_error {
  int code;
  CudaError(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};

void cuda_check(cudaError_t e, const char* what, const char* file, int line);
#define CK(x) ::kiss::cuda_check((x), #x, __FILE__, __LINE__)
#define CK_LAUNCH() ::kiss::cuda_check(cudaGetLastError(), "kernel launch", __FILE__, __LINE__)

extern thread_local u64 g_launches;  // kernels launched by this thread (reported as gpu_launches)

// ---- optional per-kernel timing (CUDA events around each launch of a kernel class, on the launching stream)
enum ProfKernel { PK_PACK, PK_LMS_CLASSIFY, PK_LMS_EMIT, PK_SCAN, PK_RADIX_HIST, PK_RADIX_SCATTER, PK_GROUPS_TINY,
                  PK_GROUPS_MEDIUM, PK_GROUPS_MEDIUM_64, PK_GROUPS_MEDIUM_128, PK_GROUPS_MEDIUM_256, PK_GROUPS_MEDIUM_512,
                  PK_GROUPS_MEDIUM_2048, PK_GROUPS_OTHER, PK_SHORTS, PK_PLACE, PK_WAVE_LMS, PK_WAVE_SCATTER, PK_WAVE_SMALL,
                  PK_COUNT };
const char* prof_name(int id);
struct Profiler {
  bool on = false;
  struct Rec { int id; cudaEvent_t a, b; };
  std::vector<Rec> recs;
  std::vector<cudaEvent_t> pool;
  double ms[PK_COUNT] = {};
  u64 launches[PK_COUNT] = {};
  u64 bytes[PK_COUNT] = {};  // algorithmic bytes of the launches of a class, where the launcher knows them
  cudaEvent_t get();
  void collect();   // call after the stream has been synchronised
  void reset();
  ~Profiler();
};
extern thread_local Profiler* g_prof;
struct ProfScope {
  Profiler* p; int id; cudaStream_t st; cudaEvent_t a;
  ProfScope(int id_, cudaStream_t st_) : p(g_prof && g_prof->on ? g_prof : nullptr), id(id_), st(st_), a(nullptr) {
    if (p) { a = p->get(); cudaEventRecord(a, st); }
  }
  ~ProfScope() {
    if (p) { cudaEvent_t b = p->get(); cudaEventRecord(b, st); p->recs.push_back({id, a, b}); }
  }
};

// ---------------------------------------------------------------- scan.cu
// in-place exclusive prefix sum of data[0..N); `partials` needs N/4096 + 2 entries of scratch;
// the grand total goes to *total_dev when non-null.
void scan_excl_u32(u32* data, u64 N, u32* partials, u32* total_dev, cudaStream_t st);
void scan_excl_u32_to(const u32* in, u32* out, u64 N, u32* partials, u32* total_dev, cudaStream_t st);
inline u64 scan_partials_len(u64 N) { return N / 4096 + 2; }

// ---------------------------------------------------------------- radix.cu
// Stable LSD radix sort of (key, val) pairs on key bits [bit_lo, bit_hi), 8 bits per pass.
// Buffers are swapped per pass; on return keys/vals point at the sorted data.
// H needs 256 * ceil(N/4096) entries, partials scan_partials_len(that).
inline u64 radix_tiles(u64 N) { return (N + 4095) / 4096; }
void radix_sort_u32(u32*& keys, u32*& vals, u32*& keys_alt, u32*& vals_alt, u32 N, int bit_lo, int bit_hi,
                    u32* H, u32* partials, cudaStream_t st);
void radix_sort_u64(u64*& keys, u32*& vals, u64*& keys_alt, u32*& vals_alt, u32 N, int bit_lo, int bit_hi,
                    u32* H, u32* partials, cudaStream_t st);
void radix_init();  // opt in to > 48 KB dynamic shared memory (once per process)

// ---------------------------------------------------------------- pack.cu
u64 packed_words(u64 n);  // words in the packed text including padding
void pack_codes(const u8* S, u64 n, u64* T, cudaStream_t st);
void ascii_to_codes(const u8* ascii, u64 n, u8* S, cudaStream_t st);
void synth_uniform(u8* S, u64 n, u64 seed, cudaStream_t st);
// *out_dev = sum_i splitmix64((i << 32) | a[i]) mod 2^64 (order-sensitive checksum of a suffix array)
void mixed_sum(const u32* a, u64 count, u64* out_dev, cudaStream_t st);
// verify.cu: bitmap = (N + 31) / 32 words, out = 4 counters (see k_verify_sa)
void verify_sa_bytes(const u8* S, u64 n, const u32* SA, u64 N, u64 k, u32 cap, u32* bitmap, u64* out, cudaStream_t st);
void verify_sa(const u64* T, u64 n, const u32* SA, u64 N, u64 k, u32 cap_words, u32* bitmap, u64* out, cudaStream_t st);

// ---------------------------------------------------------------- ingest.cu
// file bytes -> codes with the reference's read_sequence semantics (FASTA headers and newlines dropped, see ingest.cu)
u64 ingest_chunks(u64 nbytes);
void ingest(const u8* raw_dev, u64 nbytes, u8* codes_dev, u32* kept0, u32* kept1, u8* fn, u8* bfn, u8* hflag,
            u32* partials, u64* total_dev, bool fasta, cudaStream_t st);

// ---------------------------------------------------------------- lms.cu
#define LMS_CHUNK_WORDS 64ull
inline u64 lms_chunks(u64 n) { return ((n + 31) / 32 + LMS_CHUNK_WORDS - 1) / LMS_CHUNK_WORDS; }
// Classify + count: fills lmsmask (one u64 per text word, low field bits), chunk_cnt (LMS per chunk) and
// counters[8] = count[4], lms_count[4] (must be zeroed by the caller).  gp: 1 byte per chunk of scratch.
void lms_classify(const u64* T, u64 n, u8* gp, u64* lmsmask, u32* chunk_cnt, u64* counters, cudaStream_t st);
// Emit LMS positions (ascending) and their first-16-base keys from lmsmask and the exclusive chunk offsets.
void lms_emit(const u64* T, u64 n, const u64* lmsmask, const u32* chunk_off, u32* pos, u32* key, cudaStream_t st);
// *out = number of entries of the ascending list pos[0..cnt) that are <= limit
void count_le(const u32* pos, u32 cnt, u32 limit, u32* out, cudaStream_t st);

// ---------------------------------------------------------------- lms_sort.cu
struct SortParams {
  const u64* T;
  u64 n;
  u64 D;  // comparison depth in bases; KISS_UNBOUNDED = compare to the end of the text
  u32 k;  // the caller's k (only the short-suffix comparator needs it)
  // work cap of the unbounded pivot rounds (null = none): a single comparison that runs past `cap_words` 32-base words sets
  // *abort, every other comparison gives up as soon as it sees the flag, and the caller falls back to rank doubling
  u32* abort;
  u32 cap_words;
};
#define KISS_UNBOUNDED 0xFFFFFFFFFFFFFFFFull
#define TINY_MAX 8u
#define MED_MAX 2048u

struct GroupLists {  // device
  u32* glist;        // starts of groups larger than TINY_MAX
  u32 gcap;
  u32* med;          // (start, size) pairs, one list per size class: class c starts at pair med_class_off(medN, c)
  u32 medN;          // the N the class offsets are computed from (length of the sorted array)
  u32* big;          // (start, size, known common depth) triples
  u32* counters;     // [0] = #glist, [1] = (host: #med), [2] = #big, [3] = overflow flag, [4..9] = groups per medium class,
                     // [10..15] = elements per medium class
  u64* tied;         // statistics: elements found in groups
};
void groups_tiny_u32(const u32* keys, u32* pos, u32 N, SortParams sp, u64 h, GroupLists gl, cudaStream_t st);
void groups_tiny_u64(const u64* keys, u32* pos, u32 N, SortParams sp, u64 h, GroupLists gl, cudaStream_t st);
void groups_classify_u32(const u32* keys, u32 N, u32 ng, GroupLists gl, u32 h, cudaStream_t st);
// kind: 1 / 2 = pivot keys of the big path, bounded / unbounded layout (the group depth is decoded from the key)
void groups_classify_u64(const u64* keys, u32 N, u32 ng, GroupLists gl, int kind, u32 h, cudaStream_t st);
#define GL_COUNTERS 16
#define MED_CLASSES 6
// medium size classes: TINY_MAX < size <= 32, 64, 128, 256, 512, 2048.  A class holding groups larger than `lo` has at
// most N / (lo + 1) of them: its list starts at this pair index of `med`
__host__ __device__ inline u64 med_class_off(u32 N, u32 cls) {
  u64 off = 0;
  if (cls > 0) off += N / (TINY_MAX + 1) + 1;
  if (cls > 1) off += N / 33 + 1;
  if (cls > 2) off += N / 65 + 1;
  if (cls > 3) off += N / 129 + 1;
  if (cls > 4) off += N / 257 + 1;
  if (cls > 5) off += N / 513 + 1;
  return off;
}
__host__ __device__ inline u32 med_class_of(u32 size) {
  return size <= 32 ? 0u : size <= 64 ? 1u : size <= 128 ? 2u : size <= 256 ? 3u : size <= 512 ? 4u : 5u;
}
// class_groups[6] / class_elems[6]: groups and elements per size class (a class without groups is not launched)
void groups_medium(u32* pos, const u32* med, u32 medN, const u32* class_groups, const u32* class_elems, SortParams sp,
                   u64 h, cudaStream_t st);
extern u64* g_med_dbg;  // debug statistics of the medium teams (null = off)
// sizes[g] = big[3g+1]
void big_sizes(const u32* big, u32 nbig, u32* sizes, cudaStream_t st);
// worklist of the big groups (one pivot round, see lms_sort.cu): wkey = [group | side | q | bases] against the
// group's first element, wpos = suffix, wslot = slot in the main array.  Key width below the group id: 42 bits
// (bounded layout), 37 bits (unbounded layout)
void big_fill(const u32* pos_prev, const u32* slot_prev, const u32* big, const u32* boff, u32 nbig, u32 total,
              SortParams sp, bool unbounded_layout, u64* wkey, u32* wpos, u32* wslot, cudaStream_t st);
void big_writeback(u32* pos_main, const u32* wslot, const u32* wpos, u32 total, cudaStream_t st);

// key[i] = first 16 bases of suffix pos[i]
void gather_keys(const u64* T, const u32* pos, u32 N, u32* key, cudaStream_t st);

// destination rank of every key under at most 64 ascending splitters, and the per-destination counts
void dest_by_splitters(const u32* key, u32 N, const u32* spl, u32 nspl, u32* dest, u32* counts, cudaStream_t st);

// short suffixes (fewer than D bases left): exact reference comparator (kiss1_core.hpp:94-135)
// rank0: scratch of shorts_scratch_len(ns) entries
void shorts_order(const u32* shorts, u32 ns, SortParams sp, u32* rank0, u32* sorted_shorts, cudaStream_t st);
inline u64 shorts_scratch_len(u64 ns) { u64 p = 1; while (p < ns) p <<= 1; return p; }
void shorts_insertions(const u32* sorted_shorts, u32 ns, const u32* longs, u32 nl, SortParams sp, u32* ins,
                       cudaStream_t st);

// ---------------------------------------------------------------- pd.cu (rank doubling over kISS-2's words; see api.cu)
void pd_count(const u32* lms, u32 m1, u64 n, u32* cnt, cudaStream_t st);
void pd_ntail(const u32* lms, const u32* cnt, u32 m1, u64 n, u32* out, cudaStream_t st);
void pd_words(const u32* lms, const u32* first, const u32* cnt, u32 m1, const u64* T, u32 W, u32 ntail, u32* wstart,
              u32* isstart, u32* key_in, u32* idx_in, cudaStream_t st);
void pd_heads0(const u32* key, const u32* idx, u32 W, u32 ntail, u32* head, cudaStream_t st);
void pd_heads(const u64* comp, u32 A, u32* head, cudaStream_t st);
void pd_headslot(const u32* head, const u32* ord, u32 A, u32* headslot, cudaStream_t st);
void pd_rank(const u32* elem, const u32* head, const u32* ord, const u32* headslot, const u32* gslot, const u32* isnew, u32 A,
             u32* rank, cudaStream_t st);
void pd_keep(const u32* head, u32 A, u32* keep, cudaStream_t st);
void pd_compact(const u32* elem, const u32* gslot, const u32* head, const u32* keep, const u32* dst, u32 A, u32* sa_final,
                u32* elem2, u32* gslot2, u32* head2, cudaStream_t st);
void pd_key2(const u32* elem, const u32* head, const u32* ord, const u32* rank, u32 A, u32 W, u32 off, int rb, u64* comp,
             cudaStream_t st);
void pd_key2_plain(const u32* elem, const u32* rank, u32 A, u32 W, u32 off, u32* key2, cudaStream_t st);
// lists: 7 size classes of list_cap group ordinals each; counters[8] (zeroed by the caller): groups per class, [7] = groups
// > 2048 (then the round is radix-sorted instead)
void pd_seg_small(const u32* headslot, u32 ngroups, u32 A, u32* elem, u32* key2, u32* lists, u32 list_cap, u32* counters,
                  cudaStream_t st);
void pd_seg_teams(const u32* lists, u32 list_cap, const u32* cn, const u32* headslot, u32 ngroups, u32 A, u32* elem, u32* key2,
                  cudaStream_t st);
void pd_heads_seg(const u32* key2, u32 A, u32* head, u32* isnew, cudaStream_t st);
void pd_big_sizes(const u32* biglist, u32 nb, const u32* headslot, u32 ngroups, u32 A, u32* bsize, cudaStream_t st);
void pd_big_fill(const u32* biglist, const u32* boff, u32 nb, u32 B, const u32* headslot, const u32* elem, const u32* key2,
                 int rb, u64* comp, u32* val, u32* bslot, cudaStream_t st);
void pd_big_writeback(const u64* comp, const u32* val, const u32* bslot, u32 B, int rb, u32* elem, u32* key2, cudaStream_t st);
void pd_flush(const u32* elem, const u32* gslot, u32 A, u32* sa_final, cudaStream_t st);
void pd_startflags(const u32* sa_final, const u32* isstart, const u32* wstart, u32 W, u32* flag, u32* val, cudaStream_t st);
void pd_emit(const u32* val, const u32* dst, u32 W, u32* out, cudaStream_t st);

// ---------------------------------------------------------------- fmindex.cu
// N = n + 1 suffix array entries.  bwt32: ceil(N / 16) words; tot: 4 planes (stride n1 >= N / 256 + 1, a multiple of 4) of block counts (zeroed by the
// caller; scanned per plane afterwards); occ2: n2 = N / 16 + 1 rows of 4 bytes (zeroed by the caller); b32 (nullable:
// SA_INTV = 1): ceil(N / 32) words; bocc_raw: N / 64 + 1 counts (zeroed by the caller); pri: one word
// CTX (nullable): the induction's context word of every entry (low two bits = the base before the suffix when > 1)
void fm_scan(const u32* SA, const u32* CTX, u64 N, const u64* T, u32 intv, u32* bwt32, u32* tot, u64 n1, u8* occ2, u64 n2, u32* b32,
             u32* bocc_raw, u32* pri, cudaStream_t st);
void fm_occ1(const u32* tot_scanned, u64 stride, u64 n1, u32* occ1, cudaStream_t st);
void fm_sample(const u32* SA, u64 N, u32 intv, const u32* b32, const u32* bocc_scanned, u32* sa_out, cudaStream_t st);

// ---------------------------------------------------------------- induce.cu
struct Buckets {
  u64 B[5];          // bucket starts in SA; B[4] = n + 1
  u64 lms_count[4];
  u64 lms_off[4];    // exclusive prefix of lms_count
};
// K4: merge the sorted long LMS suffixes with the inserted short ones and drop them at their bucket ends
// (CTX, when non-null, receives the "context not fetched" marker for every placed suffix)
void place_lms(const u32* longs, u32 nl, const u32* sorted_shorts, const u32* ins, u32 ns, Buckets bk, u64 n,
               u32* SA, u32* CTX, cudaStream_t st);
// place an already complete order (sentinel excluded)
void place_sorted(const u32* sorted, u32 cnt, Buckets bk, u64 n, u32* SA, u32* CTX, cudaStream_t st);

struct WaveScratch {
  u32* tot_host;  // pinned host memory the kernels write directly (UVA): [0..3] totals of the wave, [4] watchdog flag
  // single-pass (decoupled look-back) waves
  u64* status;    // 4 words per tile of the largest wave, zeroed once per call
  u32* state;     // [0] tile ticket, [1] watchdog flag; zeroed once per call
  u32* tile_base; // host: tickets handed out so far
  u32* epoch;     // host: wave number
  u64* tail_host; // pinned, written by k_wave_tail: next wave start, 4 cursors, waves run
};
#ifndef WAVE_TILE
#define WAVE_TILE 8192u  // measured per 1 Gbp: 2048: 15.0 ms, 4096: 13.8 ms, 8192: 12.8 ms, 16384: 13.2 ms of waves
#endif
void wave_init();  // shared-memory attributes of the wave kernels (once per process)
inline u64 wave_tiles(u64 len) { return (len + WAVE_TILE - 1) / WAVE_TILE; }
// One induction wave over SA[lo, lo+len) (scanned downwards when `down`): every entry v > 0 whose preceding
// symbol S[v-1] is in `symmask` appends v-1 at cur[S[v-1]] (upwards) or below it (downwards).  Synchronises
// the stream and returns the per-symbol counts in tot[4].
// CTX is the per-entry context array (n + 1 words, see induce.cu).  With `lms_src` the scanned entries are
// lms_src[lo, lo+len) — the sorted LMS list itself, so the LMS suffixes are never written to SA on this path.
// `lms_ctx` (nullable, aligned like lms_src): contexts of the list entries, fetched beforehand (lms_contexts)
// `shorts`, `ins`, `ns` (with lms_src): lms_src holds only the sorted LONG suffixes and the wave reads the merged list
// (short suffix s sits at slot ins[s] + s) without it ever being materialised; lo / len index the merged list
void induce_wave(u32* SA, u32* CTX, const u64* T, const u32* lms_src, const u32* lms_ctx, u64 lo, u64 len, bool down,
                 u32 symmask, const u64 cur[4], WaveScratch ws, u32 tot[4], cudaStream_t st, const u32* shorts = nullptr,
                 const u32* ins = nullptr, u32 ns = 0);
// out[i] = context word of suffix pos[i] (the 15 bases before it, see induce.cu)
void lms_contexts(const u64* T, const u32* pos, u64 cnt, u32* out, cudaStream_t st);
// SA[0] = n (the sentinel suffix), context not fetched
void set_sentinel(u32* SA, u32* CTX, u64 n, cudaStream_t st);

// Run every remaining wave of `bucket` on the device while it fits one tile (see k_wave_tail): the wave starts
// at *p and ends at cur[bucket].  Returns the number of waves; updates *p and cur[].
#define WAVE_TAIL_LIMIT (WAVE_TILE - 8)
u32 induce_tail(u32* SA, u32* CTX, const u64* T, u32 bucket, u64* p, bool down, u32 symmask, u64 cur[4],
                WaveScratch ws, cudaStream_t st);

}  // namespace kiss
