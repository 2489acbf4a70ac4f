// api.cu — the context, the pipeline driver and the extern "C" boundary (include/kiss_b200.h).
//
// Driver = kiss1_suffix_array_dna (kiss1_core.hpp:229-268): get_lms -> lms_suffix_direct_sort_dna ->
// put_lms_suffix -> induced_sort, each stage a set of kernels on one stream.  There is no CPU path: every
// entry either runs the kernels or returns an error code.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/kiss_b200.h"
#include "kernels.h"

using namespace kiss;


static thread_local std::string g_last_error;

namespace {

// Device scratch.  One arena per context, kept between calls and sub-allocated on the host (first fit + merge),
// so the steady state makes no driver allocation at all (cudaMallocAsync pool growth showed up as 2-18 ms of
// jitter in the get_lms stage).  What does not fit goes through the stream-ordered allocator for this call and
// the arena is regrown to the observed peak once it is idle.
struct Arena {
  char* base = nullptr;
  u64 cap = 0, peak = 0, live_extra = 0;
  struct Blk { u64 off, size; bool used; };
  std::vector<Blk> blks;
  u64 top() const { return blks.empty() ? 0 : blks.back().off + blks.back().size; }
  void note() {
    u64 need = top() + live_extra;
    if (need > peak) peak = need;
  }
  void* take(u64 bytes) {
    for (size_t i = 0; i < blks.size(); i++)
      if (!blks[i].used && blks[i].size >= bytes) {
        if (blks[i].size > bytes) {
          Blk rest{blks[i].off + bytes, blks[i].size - bytes, false};
          blks[i].size = bytes;
          blks.insert(blks.begin() + i + 1, rest);
        }
        blks[i].used = true;
        return base + blks[i].off;
      }
    u64 t = top();
    if (base && t + bytes <= cap) {
      blks.push_back({t, bytes, true});
      note();
      return base + t;
    }
    return nullptr;
  }
  bool owns(void* p) const { return base && (char*)p >= base && (char*)p < base + cap; }
  void give(void* p) {
    u64 off = (u64)((char*)p - base);
    for (size_t i = 0; i < blks.size(); i++)
      if (blks[i].off == off) {
        blks[i].used = false;
        if (i + 1 < blks.size() && !blks[i + 1].used) { blks[i].size += blks[i + 1].size; blks.erase(blks.begin() + i + 1); }
        if (i > 0 && !blks[i - 1].used) { blks[i - 1].size += blks[i].size; blks.erase(blks.begin() + i); }
        while (!blks.empty() && !blks.back().used) blks.pop_back();
        return;
      }
  }
  bool idle() const { return blks.empty() && live_extra == 0; }
  // regrow to the observed peak (call only when idle and the stream is quiescent)
  void settle() {
    if (!idle() || peak <= cap) return;
    if (base) cudaFree(base);
    base = nullptr;
    cap = 0;
    u64 want = peak + peak / 16 + (64ull << 20);
    void* p = nullptr;
    if (cudaMalloc(&p, want) == cudaSuccess) { base = (char*)p; cap = want; } else cudaGetLastError();
  }
  ~Arena() { if (base) cudaFree(base); }
};

struct Pool {
  cudaStream_t st;
  Arena* ar;
  std::vector<std::pair<void*, u64>> ptrs;
  Pool(cudaStream_t s, Arena* a) : st(s), ar(a) {}
  template <typename T>
  T* alloc(u64 count) {
    u64 bytes = (count ? count : 1) * sizeof(T);
    bytes = (bytes + 255) & ~255ull;
    void* p = ar->take(bytes);
    if (!p) {
      static const bool dbg = getenv("KISS_B200_DEBUG_ARENA") != nullptr;
      if (dbg) fprintf(stderr, "[kiss_b200] arena overflow: %llu bytes (cap %llu, top %llu, blocks %zu)\n",
                       (unsigned long long)bytes, (unsigned long long)ar->cap, (unsigned long long)ar->top(), ar->blks.size());
      CK(cudaMallocAsync(&p, bytes, st));
      ar->live_extra += bytes;
      ar->note();
    }
    ptrs.push_back({p, bytes});
    return static_cast<T*>(p);
  }
  static void free_ptr(Arena* ar, cudaStream_t st, void* p, u64 bytes) {
    if (ar->owns(p)) ar->give(p);
    else { cudaFreeAsync(p, st); ar->live_extra -= bytes; }
  }
  void release(void* p) {
    for (size_t i = 0; i < ptrs.size(); i++)
      if (ptrs[i].first == p) {
        u64 bytes = ptrs[i].second;
        ptrs[i] = ptrs.back();
        ptrs.pop_back();
        free_ptr(ar, st, p, bytes);
        return;
      }
  }
  // hand ownership to the caller (who frees with Pool::free_ptr); returns the block size
  u64 keep(void* p) {
    for (size_t i = 0; i < ptrs.size(); i++)
      if (ptrs[i].first == p) {
        u64 bytes = ptrs[i].second;
        ptrs[i] = ptrs.back();
        ptrs.pop_back();
        return bytes;
      }
    return 0;
  }
  ~Pool() {
    // everything of this call is enqueued on one stream, so host-side reuse of a block by a later allocation
    // is ordered after its last use
    for (auto& pr : ptrs) free_ptr(ar, st, pr.first, pr.second);
  }
};

}  // namespace

struct kiss_b200_ctx {
  int device = 0;
  cudaStream_t st = nullptr;
  cudaEvent_t ev[8] = {};
  cudaStream_t st_copy = nullptr;  // device-to-host copies of finished SA regions, concurrent with the S pass
  cudaStream_t st_aux = nullptr;   // the tie-group teams of the LMS sort, concurrent with the big-group radix rounds
  cudaEvent_t ev_aux = nullptr;
  u32* pinned = nullptr;  // 64 u32 of pinned host memory for small read-backs (+ 8 u64 for the wave tail)
  Profiler prof;          // optional per-kernel timing
  Arena arena;
  // state kept between prepare_dev and finish_from_sorted_dev
  struct Prepared {
    bool valid = false;
    u64 n = 0, m1 = 0;      // m1 = LMS count without the sentinel
    u32 n_long = 0;
    u64* T = nullptr;
    u32* pos = nullptr;     // text-ordered LMS positions (m1)
    u32* key = nullptr;     // their first-16-base keys
    u64 bT = 0, bpos = 0, bkey = 0;  // block sizes (arena bookkeeping)
    u64 counters[8] = {};
    SortParams sp = {};
    int algo = 0;
  } prep;
};

namespace {

struct Sync {
  kiss_b200_ctx* c;
  // copy `cnt` u32 from the device and wait for them
  void read(const u32* dev, u32 cnt, u32* out) {
    CK(cudaMemcpyAsync(c->pinned, dev, cnt * sizeof(u32), cudaMemcpyDeviceToHost, c->st));
    CK(cudaStreamSynchronize(c->st));
    for (u32 i = 0; i < cnt; i++) out[i] = c->pinned[i];
  }
};

u64 depth_of(u32 k) { return 125ull * ((u64)k / 125ull + 1ull); }  // kiss1_core.hpp:94-120, constant.hpp:29

void free_prepared(kiss_b200_ctx* c) {
  auto& p = c->prep;
  if (p.T) Pool::free_ptr(&c->arena, c->st, p.T, p.bT);
  if (p.pos) Pool::free_ptr(&c->arena, c->st, p.pos, p.bpos);
  if (p.key) Pool::free_ptr(&c->arena, c->st, p.key, p.bkey);
  p = kiss_b200_ctx::Prepared();
}

// Drop the previous text and, when nothing else lives in the arena, regrow it to the peak seen so far.
// Must run before a call allocates anything (run_host allocates the device text first).
void begin_call(kiss_b200_ctx* c) {
  free_prepared(c);
  if (c->arena.idle() && c->arena.peak > c->arena.cap) {
    cudaStreamSynchronize(c->st);
    c->arena.settle();
  }
}

// ---- pack + get_lms.  Leaves T, pos, key owned by ctx->prep.
void do_prepare(kiss_b200_ctx* c, const u8* dS, u64 n, u32 k, int algo, float* t_pack, float* t_lms) {
  begin_call(c);
  cudaStream_t st = c->st;
  Pool pool(st, &c->arena);
  Sync sync{c};
  auto& P = c->prep;
  P.n = n;
  CK(cudaEventRecord(c->ev[0], st));
  u64* T = pool.alloc<u64>(packed_words(n));
  pack_codes(dS, n, T, st);
  CK(cudaEventRecord(c->ev[1], st));

  u64 nchunks = lms_chunks(n), nwords = (n + 31) / 32;
  u8* gp = pool.alloc<u8>(nchunks);
  u64* lmsmask = pool.alloc<u64>(nwords);
  u32* chunk_cnt = pool.alloc<u32>(nchunks + 1);
  u64* counters = pool.alloc<u64>(8);
  u32* partials = pool.alloc<u32>(scan_partials_len(nchunks));
  u32* total = pool.alloc<u32>(4);
  CK(cudaMemsetAsync(counters, 0, 8 * sizeof(u64), st));
  lms_classify(T, n, gp, lmsmask, chunk_cnt, counters, st);
  scan_excl_u32(chunk_cnt, nchunks, partials, total, st);
  u32 m1;
  sync.read(total, 1, &m1);
  CK(cudaMemcpyAsync(P.counters, counters, 8 * sizeof(u64), cudaMemcpyDeviceToHost, st));
  u32* pos = pool.alloc<u32>((u64)m1 + 1);
  u32* key = pool.alloc<u32>((u64)m1 + 1);
  lms_emit(T, n, lmsmask, chunk_cnt /* now exclusive offsets */, pos, key, st);

  // depth and the long / short split
  // PREFIX_DOUBLING with a bounded k: kISS-2 doubles over word indices for offset = 1, 2, 4, ... < k
  // (kiss2_core.hpp:764-766), i.e. it orders by the first 2^ceil(log2 k) words of at most 16 bases each, and leaves
  // deeper ties in an order that depends on its thread count.  Sorting to pd_depth(k) =
  // max(4096, 16 * 2^ceil(log2 k)) bases — an upper bound of that depth — and then by position is therefore equal
  // to the reference wherever the reference is tie-free, a valid k-order everywhere, and far cheaper than the full
  // order on repeat-rich text (exact 100 kb duplications: 1.02 s of LMS sort on the 3.1 Gbp chm13-like text).
  // The suffix-end rule of the unbounded comparison (the shorter first) applies: no "short" suffixes.
  u64 pd_depth = KISS_UNBOUNDED;
  if (algo == KISS_B200_PREFIX_DOUBLING && k != 0xFFFFFFFFu) {
    u64 w = 1;
    while (w < (u64)k) w <<= 1;
    pd_depth = 16ull * w < 4096ull ? 4096ull : 16ull * w;
  }
  // k >= n - 1 is the unbounded order: the lambda's depth rule (kiss1_core.hpp:131-133) needs two suffixes with >= k bases
  // each, i.e. two LMS positions <= n - k <= 1, and position 0 is never an LMS position
  bool unbounded = (algo == KISS_B200_PREFIX_DOUBLING) || ((u64)k + 1 >= n);
  P.sp.T = T; P.sp.n = n; P.sp.k = k;
  P.algo = algo;
  u32 n_long;
  if (unbounded) {
    P.sp.D = (algo == KISS_B200_PREFIX_DOUBLING && pd_depth < n) ? pd_depth : KISS_UNBOUNDED;
    n_long = m1;
    CK(cudaStreamSynchronize(st));
  } else {
    P.sp.D = depth_of(k);
    if (n >= P.sp.D && m1) {
      count_le(pos, m1, (u32)(n - P.sp.D), total, st);
      sync.read(total, 1, &n_long);
    } else {
      n_long = 0;
      CK(cudaStreamSynchronize(st));
    }
  }
  CK(cudaEventRecord(c->ev[2], st));
  CK(cudaEventSynchronize(c->ev[2]));
  if (t_pack) CK(cudaEventElapsedTime(t_pack, c->ev[0], c->ev[1]));
  if (t_lms) CK(cudaEventElapsedTime(t_lms, c->ev[1], c->ev[2]));
  P.m1 = m1; P.n_long = n_long; P.T = T; P.pos = pos; P.key = key; P.valid = true;
  P.bT = pool.keep(T); P.bpos = pool.keep(pos); P.bkey = pool.keep(key);
}

}  // namespace

namespace {

struct SortStats { u64 tied = 0, rounds = 0; };

// one level of group handling; returns the number of big groups (their list in gl.big)
template <typename K>
u32 run_groups(kiss_b200_ctx* c, const K* keys, u32* pos, u32 N, const SortParams& sp, u64 h, GroupLists gl, int kind = 0);

// The teams of a level (k_groups_medium) only touch the slots of their own groups, and so do the pivot rounds of the big
// groups (k_big_fill, the radix passes, the levels below): the teams run on a second stream next to them.  They are
// launched after a host-side wait on the main stream (the read-back of the class counts), so everything before them on
// the main stream has finished; what comes after them — the write-back of a level, the final join — is ordered on the
// second stream / by ev_aux.  KISS_B200_ONE_STREAM=1 keeps everything on the main stream (A/B timing).
cudaStream_t team_stream(kiss_b200_ctx* c) {
  static const bool one = getenv("KISS_B200_ONE_STREAM") != nullptr;
  return one ? c->st : c->st_aux;
}

// KISS_B200_DEBUG_GROUPS=1: per level, how many groups of each class and the size histogram of the medium ones
void debug_groups(kiss_b200_ctx* c, const char* tag, u32 N, u64 h, const u32* cn, const GroupLists& gl) {
  static const bool on = getenv("KISS_B200_DEBUG_GROUPS") != nullptr;
  if (!on) return;
  (void)c; (void)gl;
  fprintf(stderr, "[kiss_b200] groups %s h=%llu N=%u: >tiny %u, medium %u, big %u; medium groups (elements) of size "
                  "<=32: %u (%u) <=64: %u (%u) <=128: %u (%u) <=256: %u (%u) <=512: %u (%u) <=2048: %u (%u)\n",
          tag, (unsigned long long)h, N, cn[0], cn[1], cn[2], cn[4], cn[10], cn[5], cn[11], cn[6], cn[12], cn[7], cn[13],
          cn[8], cn[14], cn[9], cn[15]);
}

template <>
u32 run_groups<u32>(kiss_b200_ctx* c, const u32* keys, u32* pos, u32 N, const SortParams& sp, u64 h, GroupLists gl, int) {
  Sync sync{c};
  CK(cudaMemsetAsync(gl.counters, 0, GL_COUNTERS * sizeof(u32), c->st));
  groups_tiny_u32(keys, pos, N, sp, h, gl, c->st);
  u32 cn[GL_COUNTERS];
  sync.read(gl.counters, 4, cn);
  if (cn[3]) throw CudaError(-3, "internal: group list overflow");
  if (!cn[0] || h >= sp.D) return 0;
  groups_classify_u32(keys, N, cn[0], gl, (u32)h, c->st);
  sync.read(gl.counters, GL_COUNTERS, cn);
  cn[1] = cn[4] + cn[5] + cn[6] + cn[7] + cn[8] + cn[9];
  debug_groups(c, "u32", N, h, cn, gl);
  groups_medium(pos, gl.med, gl.medN, cn + 4, cn + 10, sp, h, team_stream(c));
  return cn[2];
}
template <>
u32 run_groups<u64>(kiss_b200_ctx* c, const u64* keys, u32* pos, u32 N, const SortParams& sp, u64 h, GroupLists gl, int kind) {
  Sync sync{c};
  CK(cudaMemsetAsync(gl.counters, 0, GL_COUNTERS * sizeof(u32), c->st));
  groups_tiny_u64(keys, pos, N, sp, h, gl, c->st);
  u32 cn[GL_COUNTERS];
  sync.read(gl.counters, 4, cn);
  if (cn[3]) throw CudaError(-3, "internal: group list overflow");
  if (!cn[0] || h >= sp.D) return 0;
  groups_classify_u64(keys, N, cn[0], gl, kind, (u32)h, c->st);
  sync.read(gl.counters, GL_COUNTERS, cn);
  cn[1] = cn[4] + cn[5] + cn[6] + cn[7] + cn[8] + cn[9];
  debug_groups(c, "u64", N, h, cn, gl);
  groups_medium(pos, gl.med, gl.medN, cn + 4, cn + 10, sp, h, team_stream(c));
  return cn[2];
}

GroupLists alloc_lists(Pool& pool, u32 N, u64* tied) {
  GroupLists gl;
  gl.gcap = N / (TINY_MAX + 1) + 1;
  gl.glist = pool.alloc<u32>(gl.gcap);
  gl.medN = N;
  gl.med = pool.alloc<u32>(2ull * med_class_off(N, MED_CLASSES));
  gl.big = pool.alloc<u32>(3ull * (N / (MED_MAX + 1) + 1));
  gl.counters = pool.alloc<u32>(GL_COUNTERS);
  gl.tied = tied;
  return gl;
}

// Sort `N` long LMS suffixes given in ascending text order with their first-16-base keys.
// On return *out_pos points at the sorted positions (one of pos / pos_alt).
void sort_long_lms(kiss_b200_ctx* c, Pool& pool, u32* key, u32* pos, u32 N, const SortParams& sp, u32** out_pos,
                   SortStats* stats) {
  cudaStream_t st = c->st;
  Sync sync{c};
  *out_pos = pos;
  if (N == 0) return;
  if (getenv("KISS_B200_DEBUG_GROUPS") && !g_med_dbg) {
    CK(cudaMalloc(&g_med_dbg, 6 * 64 * sizeof(u64)));
    CK(cudaMemset(g_med_dbg, 0, 6 * 64 * sizeof(u64)));
  }
  u32* key_alt = pool.alloc<u32>(N);
  u32* pos_alt = pool.alloc<u32>(N);
  u64 ntiles = radix_tiles(N);
  u32* H = pool.alloc<u32>(256 * ntiles);
  u32* partials = pool.alloc<u32>(scan_partials_len(256 * ntiles));
  radix_sort_u32(key, pos, key_alt, pos_alt, N, 0, 32, H, partials, st);
  pool.release(H);
  pool.release(partials);
  // `key` / `pos` now name the sorted buffers
  *out_pos = pos;

  u64* tied = pool.alloc<u64>(1);
  CK(cudaMemsetAsync(tied, 0, sizeof(u64), st));
  u64 h = 16;
  GroupLists gl = alloc_lists(pool, N, tied);
  u32 nbig = run_groups<u32>(c, key, pos, N, sp, h, gl);

  // pivot rounds over the big groups (lms_sort.cu): h is now only a lower bound of what a group shares
  const bool unb = sp.D == KISS_UNBOUNDED || sp.D > 4000;  // the bounded key layout holds depths < 4096
  const int field_bits = unb ? 37 : 42;
  struct Level { u64 *wkey = nullptr, *wkey_alt = nullptr; u32 *wpos = nullptr, *wpos_alt = nullptr, *wslot = nullptr; u32 total = 0; };
  const u32* pos_prev = pos;
  const u32* slot_prev = nullptr;
  GroupLists gl_prev = gl;
  while (nbig) {
    stats->rounds++;
    u32* boff = pool.alloc<u32>(nbig + 1);
    u32* bpart = pool.alloc<u32>(scan_partials_len(nbig));
    u32* btot = pool.alloc<u32>(1);
    big_sizes(gl_prev.big, nbig, boff, st);
    scan_excl_u32(boff, nbig, bpart, btot, st);
    u32 total;
    sync.read(btot, 1, &total);
    Level cur;
    cur.total = total;
    cur.wkey = pool.alloc<u64>(total); cur.wkey_alt = pool.alloc<u64>(total);
    cur.wpos = pool.alloc<u32>(total); cur.wpos_alt = pool.alloc<u32>(total);
    cur.wslot = pool.alloc<u32>(total);
    big_fill(pos_prev, slot_prev, gl_prev.big, boff, nbig, total, sp, unb, cur.wkey, cur.wpos, cur.wslot, st);
    int nb = 0;
    while (nb < 31 && (1u << nb) < nbig) nb++;
    u64 wt = radix_tiles(total);
    u32* H2 = pool.alloc<u32>(256 * wt);
    u32* part2 = pool.alloc<u32>(scan_partials_len(256 * wt));
    radix_sort_u64(cur.wkey, cur.wpos, cur.wkey_alt, cur.wpos_alt, total, 0, field_bits + nb, H2, part2, st);
    pool.release(H2); pool.release(part2); pool.release(boff); pool.release(bpart); pool.release(btot);
    h += unb ? 1 : 14;  // every class of equal keys shares at least this much
    GroupLists gl_cur = alloc_lists(pool, total, tied);
    u32 nbig_next = 0;
    if (h < sp.D) nbig_next = run_groups<u64>(c, cur.wkey, cur.wpos, total, sp, h, gl_cur, unb ? 2 : 1);
    else CK(cudaStreamSynchronize(st));  // (the write-back below runs on the team stream)
    // after this level's teams (same stream); the levels below overwrite the slots of their own groups later, in order
    big_writeback(pos, cur.wslot, cur.wpos, total, team_stream(c));
    // (nothing of a level is released before the join: its teams and its write-back may still be running)
    gl_prev = gl_cur;
    pos_prev = cur.wpos;
    slot_prev = cur.wslot;
    nbig = nbig_next;
  }
  // join the team stream
  CK(cudaEventRecord(c->ev_aux, team_stream(c)));
  CK(cudaStreamWaitEvent(st, c->ev_aux, 0));
  u64 tied_h;
  CK(cudaMemcpyAsync(&tied_h, tied, sizeof(u64), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  stats->tied = tied_h;
  if (g_med_dbg) {
    u64 d[6 * 64];
    CK(cudaMemcpy(d, g_med_dbg, sizeof(d), cudaMemcpyDeviceToHost));
    CK(cudaMemset(g_med_dbg, 0, sizeof(d)));
    for (int c = 0; c < 6; c++) {
      const u64* x = d + 64 * c;
      if (!x[0]) continue;
      fprintf(stderr, "[kiss_b200] medium class %d: %llu team-rounds (%llu network, %llu ranking); per round (groups, active elements, mean longest segment):",
              c, (unsigned long long)x[0], (unsigned long long)x[1], (unsigned long long)x[2]);
      for (int r = 0; r < 16 && x[48 + r]; r++)
        fprintf(stderr, " [%d: %llu, %llu, %.0f]", r, (unsigned long long)x[48 + r], (unsigned long long)x[16 + r],
                (double)x[32 + r] / (double)x[48 + r]);
      fprintf(stderr, "; first round by buckets %llu, by the network %llu (%llu with a crowded pivot bucket), elements equal to the pivot %llu, longest bucket %llu\n",
              (unsigned long long)x[3], (unsigned long long)x[4], (unsigned long long)x[5], (unsigned long long)x[6], (unsigned long long)x[8]);
    }
  }
}

// ---- K3' (pd.cu): rank doubling over kISS-2's word indices.  `lms` = the m1 LMS positions in text order; returns the
// sorted positions (pool-owned).  k = the caller's k: the doubling loop runs while offset < k (kiss2_core.hpp:764-766).
u32* sort_lms_doubling(kiss_b200_ctx* c, Pool& pool, const u64* T, u64 n, const u32* lms, u32 m1, u32 k, SortStats* stats) {
  cudaStream_t st = c->st;
  Sync sync{c};
  u32* out = pool.alloc<u32>(m1);
  if (!m1) return out;
  u32* cnt = pool.alloc<u32>((u64)m1 + 1);
  u32* first = pool.alloc<u32>((u64)m1 + 1);
  u32* small = pool.alloc<u32>(4);
  u32* small8 = pool.alloc<u32>(8);
  pd_count(lms, m1, n, cnt, st);
  u32* part_m = pool.alloc<u32>(scan_partials_len(m1));
  scan_excl_u32_to(cnt, first, m1, part_m, small, st);
  pd_ntail(lms, cnt, m1, n, small + 1, st);
  u32 hv[2];
  sync.read(small, 2, hv);
  pool.release(part_m);
  const u64 W64 = hv[0];
  const u32 W = hv[0], ntail = hv[1];
  if (W64 >= 0xFFFFFFF0ull) throw CudaError(KISS_B200_ERR_TOO_LARGE, "too many kISS-2 words for 32-bit indices");
  u32* wstart = pool.alloc<u32>(W);
  u32* isstart = pool.alloc<u32>(((u64)W + 31) / 32);  // bit per word: the word starts an LMS substring
  CK(cudaMemsetAsync(isstart, 0, (((u64)W + 31) / 32) * sizeof(u32), st));
  u32* key = pool.alloc<u32>(W);
  u32* idx = pool.alloc<u32>(W);
  u32* key_alt = pool.alloc<u32>(W);
  u32* idx_alt = pool.alloc<u32>(W);
  pd_words(lms, first, cnt, m1, T, W, ntail, wstart, isstart, key, idx, st);
  pool.release(cnt); pool.release(first);
  u32* H = pool.alloc<u32>(256 * radix_tiles(W));
  u32* partials = pool.alloc<u32>(scan_partials_len(256 * radix_tiles(W) > W ? 256 * radix_tiles(W) : W));
  radix_sort_u32(key, idx, key_alt, idx_alt, W, 0, 32, H, partials, st);  // stable: initialize_sa, kiss2_core.hpp:135-184
  u32* head = pool.alloc<u32>(W);
  u32* ord = pool.alloc<u32>(W);
  u32* headslot = pool.alloc<u32>(W);
  u32* rank = pool.alloc<u32>(W);
  u32* sa_final = pool.alloc<u32>(W);
  pd_heads0(key, idx, W, ntail, head, st);
  pool.release(key); pool.release(key_alt); pool.release(idx_alt);
  scan_excl_u32_to(head, ord, W, partials, nullptr, st);
  pd_headslot(head, ord, W, headslot, st);
  pd_rank(idx, head, ord, headslot, nullptr, nullptr, W, rank, st);
  // first compaction: singletons are final
  u32* keep = pool.alloc<u32>(W);
  u32* dst = pool.alloc<u32>(W);
  pd_keep(head, W, keep, st);
  scan_excl_u32_to(keep, dst, W, partials, small, st);
  u32 A;
  sync.read(small, 1, &A);
  u32* elem = pool.alloc<u32>(A);
  u32* gslot = pool.alloc<u32>(A);
  u32* ghead = pool.alloc<u32>(A);
  pd_compact(idx, nullptr, head, keep, dst, W, sa_final, elem, gslot, ghead, st);
  CK(cudaStreamSynchronize(st));
  pool.release(idx); pool.release(head); pool.release(ord); pool.release(headslot); pool.release(keep); pool.release(dst);
  pool.release(H);
  stats->tied = A;
  int rb = 1;  // bits of a rank (<= W)
  while (rb < 32 && (1ull << rb) <= (u64)W) rb++;
  for (u64 off = 1; off < (u64)k && off <= 0x7FFFFFFFull && A > 1; off *= 2) {
    stats->rounds++;
    u32* o = pool.alloc<u32>(A);
    u32* hs = pool.alloc<u32>(A);
    u64* comp = pool.alloc<u64>(A);
    u64* comp_alt = pool.alloc<u64>(A);
    u32* elem_alt = pool.alloc<u32>(A);
    u32* H2 = pool.alloc<u32>(256 * radix_tiles(A));
    scan_excl_u32_to(ghead, o, A, partials, small, st);
    u32 ngroups;
    sync.read(small, 1, &ngroups);
    // segmented sort (one pass) unless a group is larger than a block team holds
    static const bool no_seg = getenv("KISS_B200_PD_RADIX") != nullptr;  // A/B: the radix sort in every round
    bool seg_done = false;
    u32* isnew = nullptr;  // (segmented rounds) bit per slot: a head this round made
    if (!no_seg) {
      u32* key2 = reinterpret_cast<u32*>(comp);          // (the composite buffers are free on this path)
      u32* lists = reinterpret_cast<u32*>(comp_alt);   // 7 lists; a listed group has > 8 elements: at most A / 9 per list
      const u32 list_cap = A / 9 + 1;
      pd_headslot(ghead, o, A, hs, st);
      CK(cudaMemsetAsync(small8, 0, 8 * sizeof(u32), st));
      pd_key2_plain(elem, rank, A, W, (u32)off, key2, st);
      pd_seg_small(hs, ngroups, A, elem, key2, lists, list_cap, small8, st);
      u32 cn[8];
      sync.read(small8, 8, cn);
      static const bool no_mixed = getenv("KISS_B200_PD_NO_MIXED") != nullptr;  // A/B: the radix sort of everything instead
      if (cn[7] == 0 || !no_mixed) {
        pd_seg_teams(lists, list_cap, cn, hs, ngroups, A, elem, key2, st);
        if (cn[7]) {
          // the groups no block team holds: only THEIR elements go through the radix sort
          const u32 nb = cn[7];
          const u32* biglist = lists + 7ull * list_cap;
          u32* boff = pool.alloc<u32>((u64)nb + 1);
          u32* bpart = pool.alloc<u32>(scan_partials_len(nb));
          pd_big_sizes(biglist, nb, hs, ngroups, A, boff, st);
          scan_excl_u32(boff, nb, bpart, small, st);
          u32 B;
          sync.read(small, 1, &B);
          u64* bcomp = pool.alloc<u64>(B);
          u64* bcomp_alt = pool.alloc<u64>(B);
          u32* bval = pool.alloc<u32>(B);
          u32* bval_alt = pool.alloc<u32>(B);
          u32* bslot = pool.alloc<u32>(B);
          u32* Hb = pool.alloc<u32>(256 * radix_tiles(B));
          u32* partb = pool.alloc<u32>(scan_partials_len(256 * radix_tiles(B)));
          pd_big_fill(biglist, boff, nb, B, hs, elem, key2, rb, bcomp, bval, bslot, st);
          int gb = 0;
          while (gb < 31 && (1u << gb) < nb) gb++;
          radix_sort_u64(bcomp, bval, bcomp_alt, bval_alt, B, 0, rb + gb, Hb, partb, st);
          pd_big_writeback(bcomp, bval, bslot, B, rb, elem, key2, st);
          pool.release(boff); pool.release(bpart); pool.release(bcomp); pool.release(bcomp_alt); pool.release(bval);
          pool.release(bval_alt); pool.release(bslot); pool.release(Hb); pool.release(partb);
        }
        isnew = pool.alloc<u32>(((u64)A + 31) / 32);
        pd_heads_seg(key2, A, ghead, isnew, st);
        seg_done = true;
      }
    }
    if (!seg_done) {
      pd_key2(elem, ghead, o, rank, A, W, (u32)off, rb, comp, st);
      int gb = 1;
      while (gb < 32 && (1u << gb) < ngroups) gb++;
      radix_sort_u64(comp, elem, comp_alt, elem_alt, A, 0, rb + gb, H2, partials, st);
      pd_heads(comp, A, ghead, st);
    }
    scan_excl_u32_to(ghead, o, A, partials, nullptr, st);
    pd_headslot(ghead, o, A, hs, st);
    pd_rank(elem, ghead, o, hs, gslot, isnew, A, rank, st);
    u32* kp = pool.alloc<u32>(A);
    u32* ds = pool.alloc<u32>(A);
    pd_keep(ghead, A, kp, st);
    scan_excl_u32_to(kp, ds, A, partials, small, st);
    u32 A2;
    sync.read(small, 1, &A2);
    u32* elem2 = pool.alloc<u32>(A2);
    u32* gslot2 = pool.alloc<u32>(A2);
    u32* ghead2 = pool.alloc<u32>(A2);
    pd_compact(elem, gslot, ghead, kp, ds, A, sa_final, elem2, gslot2, ghead2, st);
    CK(cudaStreamSynchronize(st));
    pool.release(o); pool.release(hs); pool.release(comp); pool.release(comp_alt); pool.release(elem); pool.release(elem_alt);
    pool.release(H2); pool.release(kp); pool.release(ds); pool.release(gslot); pool.release(ghead);
    if (isnew) pool.release(isnew);
    elem = elem2; gslot = gslot2; ghead = ghead2;
    A = A2;
  }
  pd_flush(elem, gslot, A, sa_final, st);  // bounded k: what is still tied keeps ascending text position
  u32* flag = pool.alloc<u32>(W);
  pd_startflags(sa_final, isstart, wstart, W, flag, rank, st);  // (the rank array is free now: it takes the positions)
  scan_excl_u32(flag, W, partials, nullptr, st);
  pd_emit(rank, flag, W, out, st);
  CK(cudaStreamSynchronize(st));
  pool.release(flag); pool.release(elem); pool.release(gslot); pool.release(ghead); pool.release(rank); pool.release(sa_final);
  pool.release(wstart); pool.release(isstart); pool.release(partials); pool.release(small); pool.release(small8);
  return out;
}

Buckets make_buckets(const u64 counters[8], u64 n) {
  Buckets bk;
  u64 acc = 1;  // slot 0 holds the sentinel (kiss_common.hpp:463: scan with initial value 1)
  for (int s = 0; s < 4; s++) {
    bk.B[s] = acc;
    acc += counters[s];
  }
  bk.B[4] = acc;  // == n + 1
  (void)n;
  u64 off = 0;
  for (int s = 0; s < 4; s++) {
    bk.lms_count[s] = counters[4 + s];
    bk.lms_off[s] = off;
    off += counters[4 + s];
  }
  return bk;
}

struct InduceStats { u64 waves_L = 0, waves_S = 0; };

// Where finished parts of the suffix array go while induction is still running: the L-type part of every bucket
// is final after the L pass, the S-type part of bucket c after its S pass, and nothing later writes them — so
// their device-to-host copies run on a second stream under the remaining waves (a pinned destination only:
// a pageable one would block the host thread that drives the waves).
struct CopyOut {
  u32* host = nullptr;
  cudaStream_t st = nullptr;
  u64 bytes = 0;
  void send(const u32* SA, u64 lo, u64 cnt) {
    if (!host || !cnt) return;
    CK(cudaMemcpyAsync(host + lo, SA + lo, cnt * sizeof(u32), cudaMemcpyDeviceToHost, st));
    bytes += cnt * sizeof(u32);
  }
};

// `lms` = the sorted LMS suffixes (sentinel excluded), grouped by first symbol: bk.lms_off / bk.lms_count
void do_induce(kiss_b200_ctx* c, Pool& pool, const u64* T, u64 n, const Buckets& bk, const u32* lms, u32* SA, u32* CTX,
               InduceStats* is, CopyOut* co = nullptr, const u32* lms_ctx = nullptr, const u32* shorts = nullptr,
               const u32* ins = nullptr, u32 ns = 0) {
  cudaStream_t st = c->st;
  WaveScratch ws;
  u64 maxlen = n + 1;
  u64 nt = wave_tiles(maxlen);
  ws.tot_host = c->pinned + 32;
  ws.tot_host[4] = 0;
  ws.tail_host = reinterpret_cast<u64*>(c->pinned + 64);
  ws.status = pool.alloc<u64>(4 * nt + 4);
  ws.state = pool.alloc<u32>(2);
  u32 tile_base = 0, epoch = 0;
  ws.tile_base = &tile_base;
  ws.epoch = &epoch;
  CK(cudaMemsetAsync(ws.status, 0, (4 * nt + 4) * sizeof(u64), st));
  CK(cudaMemsetAsync(ws.state, 0, 2 * sizeof(u32), st));
  u32 tot[4];
  u64 Lc[4], nL[4], Sc[4];
  for (int s = 0; s < 4; s++) Lc[s] = bk.B[s];
  // L pass (kiss_common.hpp:372-403)
  set_sentinel(SA, CTX, n, st);
  induce_wave(SA, CTX, T, nullptr, nullptr, 0, 1, false, 0xFu, Lc, ws, tot, st);
  for (int s = 0; s < 4; s++) Lc[s] += tot[s];
  is->waves_L++;
  if (co) co->send(SA, 0, 1);
  for (int cb = 0; cb < 4; cb++) {
    u64 p = bk.B[cb];
    while (p < Lc[cb]) {
      u64 q = Lc[cb];
      if (q - p <= WAVE_TAIL_LIMIT) {
        is->waves_L += induce_tail(SA, CTX, T, (u32)cb, &p, false, (0xFu << cb) & 0xFu, Lc, ws, st);
        continue;
      }
      induce_wave(SA, CTX, T, nullptr, nullptr, p, q - p, false, (0xFu << cb) & 0xFu, Lc, ws, tot, st);
      for (int s = 0; s < 4; s++) Lc[s] += tot[s];
      p = q;
      is->waves_L++;
    }
    // bucket cb's L-type part is final (every wave ended with a stream synchronisation; later waves only fill
    // higher buckets): it can leave now
    if (co) co->send(SA, bk.B[cb], Lc[cb] - bk.B[cb]);
    if (bk.lms_count[cb]) {
      induce_wave(SA, CTX, T, lms, lms_ctx, bk.lms_off[cb], bk.lms_count[cb], false, 0xFu, Lc, ws, tot, st, shorts, ins, ns);
      for (int s = 0; s < 4; s++) Lc[s] += tot[s];
      is->waves_L++;
    }
  }
  for (int s = 0; s < 4; s++) {
    nL[s] = Lc[s] - bk.B[s];
    Sc[s] = bk.B[s + 1];
  }

  // S pass (kiss_common.hpp:192-224); the LMS slots are simply overwritten (induced_clear, :405-420)
  for (int cb = 3; cb >= 0; cb--) {
    u64 p = bk.B[cb + 1];
    while (Sc[cb] < p) {
      u64 q = Sc[cb];
      if (p - q <= WAVE_TAIL_LIMIT) {
        is->waves_S += induce_tail(SA, CTX, T, (u32)cb, &p, true, (1u << (cb + 1)) - 1u, Sc, ws, st);
        continue;
      }
      induce_wave(SA, CTX, T, nullptr, nullptr, q, p - q, true, (1u << (cb + 1)) - 1u, Sc, ws, tot, st);
      for (int s = 0; s < 4; s++) Sc[s] -= tot[s];
      p = q;
      is->waves_S++;
    }
    if (co) co->send(SA, bk.B[cb] + nL[cb], bk.B[cb + 1] - bk.B[cb] - nL[cb]);  // bucket cb's S-type part is final
    if (cb > 0 && nL[cb]) {
      induce_wave(SA, CTX, T, nullptr, nullptr, bk.B[cb], nL[cb], true, (1u << cb) - 1u, Sc, ws, tot, st);
      for (int s = 0; s < 4; s++) Sc[s] -= tot[s];
      is->waves_S++;
    }
  }
  if (ws.tot_host[4]) throw CudaError(-3, "internal: induction look-back watchdog fired");
  for (int s = 0; s < 4; s++)
    if (Sc[s] != bk.B[s] + nL[s]) throw CudaError(-3, "internal: induction did not fill every bucket");
}

// lms sort of the prepared text: sorted long suffixes + ordered shorts + their insertion points
struct SortedLms {
  u32* longs = nullptr; u32 nl = 0;
  u32* shorts = nullptr; u32* ins = nullptr; u32 ns = 0;
};

void do_shorts(kiss_b200_ctx* c, Pool& pool, const u32* shorts_text, u32 ns, const u32* longs, u32 nl,
               const SortParams& sp, SortedLms* out) {
  out->ns = ns;
  if (!ns) return;
  u32* rank0 = pool.alloc<u32>(shorts_scratch_len(ns));
  out->shorts = pool.alloc<u32>(ns);
  out->ins = pool.alloc<u32>(ns);
  shorts_order(shorts_text, ns, sp, rank0, out->shorts, c->st);
  shorts_insertions(out->shorts, ns, longs, nl, sp, out->ins, c->st);
}

void do_sort(kiss_b200_ctx* c, Pool& pool, SortedLms* out, SortStats* stats) {
  auto& P = c->prep;
  u32 ns = (u32)(P.m1 - P.n_long);
  // PREFIX_DOUBLING, and the unbounded order with either -s, go through rank doubling (pd.cu): O(log) rounds whatever
  // the repeats look like.  KISS_B200_PD=pivot keeps the pivot rounds of lms_sort.cu at the depth of do_prepare (A/B timing;
  // no bound on texts with very long exact duplicates).  Tiny texts (n < 64) take the pivot path: the reference's own
  // tail-word rule degenerates there (its `n - l` wraps for n < 16, kiss2_core.hpp:81).
  static const bool pd_pivot = [] { const char* e = getenv("KISS_B200_PD"); return e && std::string(e) == "pivot"; }();
  const bool unbounded = P.sp.D == KISS_UNBOUNDED;
  if (!pd_pivot && P.n >= 64 && ns == 0 && P.m1 && (P.algo == KISS_B200_PREFIX_DOUBLING || unbounded)) {
    const u32 kk = P.algo == KISS_B200_PREFIX_DOUBLING ? P.sp.k : 0xFFFFFFFFu;
    // The unbounded order first tries the pivot rounds under a work cap: they skip a common prefix in one step where
    // doubling needs log2 of it (C4, 512 near-identical copies: 0.083 s against 0.26 s), but cost O(LCP / 32) loads per
    // element.  A comparison that runs past 16384 bases raises the abort flag, everything in flight gives up within 32
    // words, and the sort starts over as rank doubling (two exact 100 Mbp copies: aborted after a few ms).
    // KISS_B200_PD=doubling skips the attempt.
    static const bool pd_only = [] { const char* e = getenv("KISS_B200_PD"); return e && std::string(e) == "doubling"; }();
    if (kk == 0xFFFFFFFFu && !pd_only) {
      u32* flag = pool.alloc<u32>(1);
      CK(cudaMemsetAsync(flag, 0, sizeof(u32), c->st));
      SortParams sp = P.sp;
      sp.D = KISS_UNBOUNDED;
      sp.abort = flag;
      sp.cap_words = 512;
      u32* sorted;
      SortStats ss;
      sort_long_lms(c, pool, P.key, P.pos, (u32)P.m1, sp, &sorted, &ss);
      u32 aborted = 0;
      CK(cudaMemcpyAsync(&aborted, flag, sizeof(u32), cudaMemcpyDeviceToHost, c->st));
      CK(cudaStreamSynchronize(c->st));
      if (!aborted) {
        *stats = ss;
        out->longs = sorted;
        out->nl = (u32)P.m1;
        out->ns = 0;
        return;
      }
      // back to text order (ascending positions) for the doubling: P.pos / its twin hold a permutation of the list
      u32* vals = pool.alloc<u32>(P.m1);
      u32* keys_alt = pool.alloc<u32>(P.m1);
      u32* vals_alt = pool.alloc<u32>(P.m1);
      u32* H = pool.alloc<u32>(256 * radix_tiles(P.m1));
      u32* partials = pool.alloc<u32>(scan_partials_len(256 * radix_tiles(P.m1)));
      u32* keys = sorted;
      radix_sort_u32(keys, vals, keys_alt, vals_alt, (u32)P.m1, 0, 32, H, partials, c->st);
      if (keys != P.pos) CK(cudaMemcpyAsync(P.pos, keys, P.m1 * sizeof(u32), cudaMemcpyDeviceToDevice, c->st));
      CK(cudaStreamSynchronize(c->st));
      pool.release(vals); pool.release(vals_alt); pool.release(H); pool.release(partials); pool.release(flag);
      stats->rounds = 1000;  // (marks the fallback in the timings: rounds >= 1000)
    }
    SortStats sd;
    out->longs = sort_lms_doubling(c, pool, P.T, P.n, P.pos, (u32)P.m1, kk, &sd);
    stats->tied = sd.tied;
    stats->rounds += sd.rounds;
    out->nl = (u32)P.m1;
    out->ns = 0;
    return;
  }
  // the shorts are the tail of the text-ordered list; copy them out before the sort permutes the buffers
  u32* shorts_text = pool.alloc<u32>(ns);
  if (ns) CK(cudaMemcpyAsync(shorts_text, P.pos + P.n_long, ns * sizeof(u32), cudaMemcpyDeviceToDevice, c->st));
  u32* sorted;
  sort_long_lms(c, pool, P.key, P.pos, P.n_long, P.sp, &sorted, stats);
  out->longs = sorted;
  out->nl = P.n_long;
  do_shorts(c, pool, shorts_text, ns, sorted, P.n_long, P.sp, out);
}

void fill_timings(kiss_b200_timings* t, float pack, float lms, float sort, float put, float induce) {
  t->pack = pack * 1e-3; t->get_lms = lms * 1e-3; t->lms_sort = sort * 1e-3; t->put_lms = put * 1e-3;
  t->induce = induce * 1e-3;
  t->total_device = t->pack + t->get_lms + t->lms_sort + t->put_lms + t->induce;
}

enum Stop { STOP_LMS, STOP_SORT, STOP_PUT, STOP_FULL };

// the whole path on device-resident data; `stage_out` receives the stage's array when stopping early
// `keep_ctx` (n + 1 words, caller-owned): the induction's context array is left there (the FM-index scan reads it)
void run_pipeline(kiss_b200_ctx* c, const u8* dS, u64 n, u32 k, int algo, u32* dSA, Stop stop, kiss_b200_timings* t,
                  CopyOut* co = nullptr, u32* keep_ctx = nullptr) {
  cudaStream_t st = c->st;
  u64 launches0 = g_launches;
  float t_pack = 0, t_lms = 0, t_sort = 0, t_put = 0, t_ind = 0;
  do_prepare(c, dS, n, k, algo, &t_pack, &t_lms);
  auto& P = c->prep;
  SortStats ss;
  InduceStats is;
  if (stop != STOP_LMS) {
    Pool pool(st, &c->arena);
    SortedLms sl;
    u32* dCTX = nullptr;
    const u32* lms = nullptr;
    CK(cudaEventRecord(c->ev[3], st));
    do_sort(c, pool, &sl, &ss);
    CK(cudaEventRecord(c->ev[4], st));
    Buckets bk = make_buckets(P.counters, n);
    if (stop == STOP_SORT) {
      // merged list, sentinel first: "place" into a linear array
      Buckets lin = bk;
      lin.B[1] = P.m1 + 1; lin.lms_count[0] = P.m1; lin.lms_off[0] = 0;
      lin.lms_off[1] = lin.lms_off[2] = lin.lms_off[3] = ~0ull;
      place_lms(sl.longs, sl.nl, sl.shorts, sl.ins, sl.ns, lin, n, dSA, nullptr, st);
    } else if (stop == STOP_PUT) {
      CK(cudaMemsetAsync(dSA, 0xFF, (n + 1) * sizeof(u32), st));
      place_lms(sl.longs, sl.nl, sl.shorts, sl.ins, sl.ns, bk, n, dSA, nullptr, st);
    } else {
      // the full path never places: the L pass scans the sorted list itself (put_lms_suffix is folded into it)
      dCTX = keep_ctx ? keep_ctx : pool.alloc<u32>(n + 1);
      lms = sl.longs;  // the LMS waves read the merged list on the fly (induce_wave, VIRT): no merge pass
    }
    CK(cudaEventRecord(c->ev[5], st));
    if (stop == STOP_FULL) do_induce(c, pool, P.T, n, bk, lms, dSA, dCTX, &is, co, nullptr, sl.shorts, sl.ins, sl.ns);
    CK(cudaEventRecord(c->ev[6], st));
    CK(cudaEventSynchronize(c->ev[6]));
    CK(cudaEventElapsedTime(&t_sort, c->ev[3], c->ev[4]));
    CK(cudaEventElapsedTime(&t_put, c->ev[4], c->ev[5]));
    CK(cudaEventElapsedTime(&t_ind, c->ev[5], c->ev[6]));
  }
  if (t) {
    fill_timings(t, t_pack, t_lms, t_sort, t_put, t_ind);
    t->n = n; t->m = P.m1 + 1;
    t->lms_sort_tied = ss.tied; t->lms_sort_rounds = ss.rounds;
    t->waves_L = is.waves_L; t->waves_S = is.waves_S;
    t->kernel_launches = g_launches - launches0;
  }
}

int check_n(u64 n) {
  if (n + 1 >= 0xFFFFFFFFull) {
    g_last_error = "text too long for 32-bit suffix indices (EMPTY = 0xFFFFFFFF is reserved)";
    return KISS_B200_ERR_TOO_LARGE;
  }
  return 0;
}

template <typename F>
int guarded(kiss_b200_ctx* c, F&& f) {
  if (!c) { g_last_error = "null context"; return KISS_B200_ERR_ARG; }
  try {
    CK(cudaSetDevice(c->device));
    g_prof = &c->prof;
    f();
    cudaStreamSynchronize(c->st);
    cudaStreamSynchronize(c->st_copy);
    cudaStreamSynchronize(c->st_aux);
    c->prof.collect();
    c->arena.settle();
    return KISS_B200_OK;
  } catch (const CudaError& e) {
    g_last_error = e.what();
    cudaStreamSynchronize(c->st_copy);  // no copy into the caller's buffer may outlive the call
    cudaStreamSynchronize(c->st_aux);
    cudaGetLastError();
    return e.code;
  } catch (const std::bad_alloc&) {
    g_last_error = "host allocation failed";
    return KISS_B200_ERR_NOMEM;
  } catch (const std::exception& e) {
    g_last_error = e.what();
    return KISS_B200_ERR_CUDA;
  }
}

double elapsed_s(kiss_b200_ctx* c, int a, int b) {
  float ms = 0;
  cudaEventElapsedTime(&ms, c->ev[a], c->ev[b]);
  return ms * 1e-3;
}

// host text -> device, run, device SA -> host
enum Input { IN_CODES = 0, IN_ASCII = 1, IN_FILE = 2 };

// file bytes on the device -> codes on the device (ingest.cu); returns n
u64 do_ingest(kiss_b200_ctx* c, Pool& pool, const u8* dRaw, u64 nbytes, bool fasta, u8* dCodes) {
  cudaStream_t st = c->st;
  const u64 nch = ingest_chunks(nbytes);
  if (!nch) return 0;
  u32* kept0 = pool.alloc<u32>(nch);
  u32* kept1 = pool.alloc<u32>(nch);
  u8* fn = pool.alloc<u8>(nch);
  u8* bfn = pool.alloc<u8>(nch / 1024 + 2);
  u8* hflag = pool.alloc<u8>(nch);
  u32* partials = pool.alloc<u32>(scan_partials_len(nch));
  u64* total = pool.alloc<u64>(1);
  ingest(dRaw, nbytes, dCodes, kept0, kept1, fn, bfn, hflag, partials, total, fasta, st);
  u64 n = 0;
  CK(cudaMemcpyAsync(&n, total, sizeof(u64), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  pool.release(kept0); pool.release(kept1); pool.release(fn); pool.release(bfn); pool.release(hflag);
  pool.release(partials);
  pool.release(total);
  return n;
}

// `input`: what S_host holds — codes 0..3, the bare sequence in ASCII, or the bytes of a FASTA / text file (then n is
// the file size, the text length goes to *n_text and out_host must hold out_count_hint >= text length + 1 entries)
void run_host(kiss_b200_ctx* c, const u8* S_host, u64 n, u32 k, int algo, u32* out_host, u64 out_count_hint, Stop stop,
              int input, kiss_b200_timings* t, uint64_t* m_out, uint64_t* n_text = nullptr) {
  cudaStream_t st = c->st;
  begin_call(c);
  Pool pool(st, &c->arena);
  u8* dS = pool.alloc<u8>(n + 16);
  struct Events {  // (destroyed on every exit path, exceptions included)
    cudaEvent_t e[4] = {};
    Events() { for (auto& x : e) CK(cudaEventCreate(&x)); }
    ~Events() { for (auto& x : e) if (x) cudaEventDestroy(x); }
  } evs;
  cudaEvent_t e0 = evs.e[0], e1 = evs.e[1], e2 = evs.e[2], e3 = evs.e[3];
  CK(cudaEventRecord(e0, st));
  if (n) CK(cudaMemcpyAsync(dS, S_host, n, cudaMemcpyHostToDevice, st));
  if (input == IN_ASCII) {
    u8* dC = pool.alloc<u8>(n + 16);
    ascii_to_codes(dS, n, dC, st);
    dS = dC;
  } else if (input == IN_FILE) {
    u8* dC = pool.alloc<u8>(n + 16);
    const bool fasta = n && S_host[0] == '>';  // io.hpp:8
    const u64 nt = do_ingest(c, pool, dS, n, fasta, dC);
    pool.release(dS);
    dS = dC;
    n = nt;
    if (n_text) *n_text = n;
    if (n + 1 >= 0xFFFFFFFFull) throw CudaError(KISS_B200_ERR_TOO_LARGE, "text too large for 32-bit suffix indices");
    if (n + 1 > out_count_hint) throw CudaError(KISS_B200_ERR_ARG, "suffix array buffer smaller than text length + 1");
    if (n == 0) {  // headers / empty lines only: SA = {0} (kiss1_core.hpp:237-238)
      if (out_host) out_host[0] = 0;
      if (t) { memset(t, 0, sizeof(*t)); t->m = 1; }  // (as trivial_n0: the sentinel is the only LMS entry)
      if (m_out) *m_out = 1;
      return;
    }
  }
  CK(cudaEventRecord(e1, st));
  u64 out_n = (stop == STOP_LMS || stop == STOP_SORT) ? n / 2 + 2 : n + 1;
  u32* dOut = pool.alloc<u32>(out_n);
  kiss_b200_timings tt;
  memset(&tt, 0, sizeof(tt));
  CopyOut co;
  if (stop == STOP_FULL && out_host && n) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, out_host) == cudaSuccess && at.type == cudaMemoryTypeHost) {
      co.host = out_host;
      co.st = c->st_copy;
    } else {
      cudaGetLastError();
    }
  }
  run_pipeline(c, dS, n, k, algo, dOut, stop, &tt, co.host ? &co : nullptr);
  u64 m = c->prep.m1 + 1;
  u64 cnt = out_n;
  if (stop == STOP_LMS) {
    // text-ordered list + sentinel
    cnt = m;
    if (c->prep.m1) CK(cudaMemcpyAsync(dOut, c->prep.pos, c->prep.m1 * sizeof(u32), cudaMemcpyDeviceToDevice, st));
    u32 nn = (u32)n;
    CK(cudaMemcpyAsync(dOut + c->prep.m1, &nn, sizeof(u32), cudaMemcpyHostToDevice, st));
    CK(cudaStreamSynchronize(st));
  } else if (stop == STOP_SORT) {
    cnt = m;
  }
  (void)out_count_hint;
  CK(cudaEventRecord(e2, st));
  if (co.host) {
    // everything was sent region by region; what is left of the copies is the exposed device-to-host time
    if (co.bytes != cnt * sizeof(u32)) throw CudaError(-3, "internal: overlapped copy-out missed a region");
    CK(cudaEventRecord(e3, co.st));
  } else {
    if (out_host) CK(cudaMemcpyAsync(out_host, dOut, cnt * sizeof(u32), cudaMemcpyDeviceToHost, st));
    CK(cudaEventRecord(e3, st));
  }
  CK(cudaEventSynchronize(e3));
  float h2d = 0, d2h = 0;
  cudaEventElapsedTime(&h2d, e0, e1);
  cudaEventElapsedTime(&d2h, e2, e3);
  tt.h2d = h2d * 1e-3; tt.d2h = d2h * 1e-3;
  if (t) *t = tt;
  if (m_out) *m_out = m;
}

// ---- FM-index image (fmindex.cu): the bytes FMIndex<SA_INTV, uint32_t, Sorter>{.LOOKUP_LEN = 0}::save writes
// (fm_index.hpp:591-617; Serializer::save, serializer.hpp:95-112: element count as size_t, then the raw bytes)
struct FmLayout {
  u64 N, n1, n2, nb, nsa, b_words;
  u64 o_bwt, o_occ1, o_occ2, o_sa, o_lookup, o_b, o_bocc, total;  // offsets of the DATA of each section (its count sits 8 bytes before)
  u64 by_bwt, by_occ1, by_occ2, by_sa, by_b, by_bocc;
};
FmLayout fm_layout(u64 n, u32 intv) {
  FmLayout L{};
  L.N = n + 1;
  L.n1 = L.N / 256 + 1; L.n2 = L.N / 16 + 1; L.nb = L.N / 64 + 1;
  L.nsa = intv == 1 ? L.N : (L.N + intv - 1) / intv;
  L.b_words = (L.N + 63) / 64;
  L.by_bwt = (L.N + 3) / 4; L.by_occ1 = 16 * L.n1; L.by_occ2 = 4 * L.n2; L.by_sa = 4 * L.nsa;
  L.by_b = 8 * L.b_words; L.by_bocc = 4 * L.nb;
  u64 off = 20;  // cnt_[4], pri_
  L.o_bwt = off + 8; off = L.o_bwt + L.by_bwt;
  L.o_occ1 = off + 8; off = L.o_occ1 + L.by_occ1;
  L.o_occ2 = off + 8; off = L.o_occ2 + L.by_occ2;
  L.o_sa = off + 8; off = L.o_sa + L.by_sa;
  L.o_lookup = off + 8; off = L.o_lookup + 8;
  if (intv != 1) {
    L.o_b = off + 8; off = L.o_b + L.by_b;
    L.o_bocc = off + 8; off = L.o_bocc + L.by_bocc;
  }
  L.total = off;
  return L;
}

void put_u64(u8* p, u64 v) { memcpy(p, &v, 8); }

// T = packed text (may be null when n == 0), dSA = the n + 1 suffix array entries on the device; writes the image
void do_fmindex(kiss_b200_ctx* c, Pool& pool, const u64* T, u64 n, const u32* dSA, const u32* dCTX, u32 intv, u8* image,
                float* ms) {
  cudaStream_t st = c->st;
  const FmLayout L = fm_layout(n, intv);
  CK(cudaEventRecord(c->ev[7], st));
  u32* bwt32 = pool.alloc<u32>((L.N + 15) / 16 + 4);
  const u64 n1p = (L.n1 + 3) & ~3ull;  // plane stride: the scan reads 16 bytes at a time
  u32* tot = pool.alloc<u32>(4 * n1p);
  u8* occ2 = pool.alloc<u8>(4 * L.n2);
  u32* occ1 = pool.alloc<u32>(4 * L.n1);
  u32* totals = pool.alloc<u32>(8);  // [0..3] symbol totals, [4] pri
  u32* partials = pool.alloc<u32>(scan_partials_len(L.n1 > L.nb ? L.n1 : L.nb));
  u32 *b32 = nullptr, *bocc = nullptr, *sa_out = nullptr;
  CK(cudaMemsetAsync(tot, 0, 4 * n1p * sizeof(u32), st));
  CK(cudaMemsetAsync(occ2, 0, 4 * L.n2, st));
  if (intv != 1) {
    b32 = pool.alloc<u32>(2 * L.b_words);
    bocc = pool.alloc<u32>(L.nb);
    sa_out = pool.alloc<u32>(L.nsa);
    CK(cudaMemsetAsync(b32 + 2 * (L.b_words - 1), 0, 8, st));  // the last word may be written by half
    CK(cudaMemsetAsync(bocc, 0, L.nb * sizeof(u32), st));
  }
  fm_scan(dSA, dCTX, L.N, T, intv, bwt32, tot, n1p, occ2, L.n2, b32, bocc, totals + 4, st);
  // what the scan finished leaves on the copy stream while the prefix sums and the compaction run
  CK(cudaEventRecord(c->ev_aux, st));
  CK(cudaStreamWaitEvent(c->st_copy, c->ev_aux, 0));
  CK(cudaMemcpyAsync(image + L.o_bwt, bwt32, L.by_bwt, cudaMemcpyDeviceToHost, c->st_copy));
  CK(cudaMemcpyAsync(image + L.o_occ2, occ2, L.by_occ2, cudaMemcpyDeviceToHost, c->st_copy));
  if (intv != 1) CK(cudaMemcpyAsync(image + L.o_b, b32, L.by_b, cudaMemcpyDeviceToHost, c->st_copy));
  for (int s = 0; s < 4; s++) scan_excl_u32(tot + s * n1p, L.n1, partials, totals + s, st);
  fm_occ1(tot, n1p, L.n1, occ1, st);
  if (intv != 1) {
    scan_excl_u32(bocc, L.nb, partials, nullptr, st);
    fm_sample(dSA, L.N, intv, b32, bocc, sa_out, st);
  }
  CK(cudaEventRecord(c->ev[0], st));
  u32 h[5];
  CK(cudaMemcpyAsync(h, totals, 5 * sizeof(u32), cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(image + L.o_occ1, occ1, L.by_occ1, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(image + L.o_sa, intv == 1 ? dSA : sa_out, L.by_sa, cudaMemcpyDeviceToHost, st));
  if (intv != 1) CK(cudaMemcpyAsync(image + L.o_bocc, bocc, L.by_bocc, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  CK(cudaStreamSynchronize(c->st_copy));
  // cnt_[c] = 1 + number of symbols smaller than c (fm_index.hpp:305-309); pri_; the section counts; lookup_ = {0, N}
  u32 cnt[4], acc = 1;
  for (int s = 0; s < 4; s++) { cnt[s] = acc; acc += h[s]; }
  memcpy(image, cnt, 16);
  memcpy(image + 16, &h[4], 4);
  put_u64(image + L.o_bwt - 8, L.N);
  put_u64(image + L.o_occ1 - 8, L.n1);
  put_u64(image + L.o_occ2 - 8, L.n2);
  put_u64(image + L.o_sa - 8, L.nsa);
  put_u64(image + L.o_lookup - 8, 2);
  const u32 lookup[2] = {0u, (u32)L.N};
  memcpy(image + L.o_lookup, lookup, 8);
  if (intv != 1) {
    put_u64(image + L.o_b - 8, L.N);
    put_u64(image + L.o_bocc - 8, L.nb);
  }
  if (ms) CK(cudaEventElapsedTime(ms, c->ev[7], c->ev[0]));
}

// host text (codes or file bytes) -> suffix array on the device -> FM-index image on the host
void run_fm_host(kiss_b200_ctx* c, const u8* S_host, u64 n, int input, u32 k, int algo, u32 intv, u8* image, u64 capacity,
                 uint64_t* image_bytes, uint64_t* n_text, kiss_b200_timings* t, double* fm_seconds) {
  cudaStream_t st = c->st;
  begin_call(c);
  Pool pool(st, &c->arena);
  u8* dS = pool.alloc<u8>(n + 16);
  const auto w0 = std::chrono::steady_clock::now();
  if (n) CK(cudaMemcpyAsync(dS, S_host, n, cudaMemcpyHostToDevice, st));
  if (input == IN_FILE) {
    u8* dC = pool.alloc<u8>(n + 16);
    const bool fasta = n && S_host[0] == '>';
    n = do_ingest(c, pool, dS, n, fasta, dC);
    pool.release(dS);
    dS = dC;
    if (n + 1 >= 0xFFFFFFFFull) throw CudaError(KISS_B200_ERR_TOO_LARGE, "text too large for 32-bit suffix indices");
  }
  if (n_text) *n_text = n;
  const FmLayout L = fm_layout(n, intv);
  *image_bytes = L.total;
  if (L.total > capacity) throw CudaError(KISS_B200_ERR_ARG, "image buffer smaller than kiss_b200_fmindex_image_bytes(n, sa_intv)");
  CK(cudaStreamSynchronize(st));
  const double h2d = std::chrono::duration<double>(std::chrono::steady_clock::now() - w0).count();
  u32* dSA = pool.alloc<u32>(n + 1);
  u32* dCTX = n ? pool.alloc<u32>(n + 1) : nullptr;
  kiss_b200_timings tt;
  memset(&tt, 0, sizeof(tt));
  const u64* T = nullptr;
  if (n) {
    run_pipeline(c, dS, n, k, algo, dSA, STOP_FULL, &tt, nullptr, dCTX);
    T = c->prep.T;
  } else {
    CK(cudaMemsetAsync(dSA, 0, sizeof(u32), st));
    tt.m = 1;
  }
  pool.release(dS);
  float ms = 0;
  const auto w1 = std::chrono::steady_clock::now();
  do_fmindex(c, pool, T, n, dSA, dCTX, intv, image, &ms);
  tt.h2d = h2d;
  tt.d2h = std::chrono::duration<double>(std::chrono::steady_clock::now() - w1).count() - ms * 1e-3;
  if (t) *t = tt;
  if (fm_seconds) *fm_seconds = ms * 1e-3;
  free_prepared(c);
}

}  // namespace

// ---- accessors for the entry points that live in other files (noalloc.cu)
namespace kiss {
void api_set_last_error(const std::string& msg) { g_last_error = msg; }
int api_ctx_device(const kiss_b200_ctx* c) { return c->device; }
}  // namespace kiss

// ======================================================================================= extern "C"
extern "C" {

int kiss_b200_abi_version(void) { return 1; }

const char* kiss_b200_strerror(int code) {
  switch (code) {
    case KISS_B200_OK: return "ok";
    case KISS_B200_ERR_ARG: return "invalid argument";
    case KISS_B200_ERR_TOO_LARGE: return "text too large for 32-bit suffix indices";
    case KISS_B200_ERR_CUDA: return "CUDA error";
    case KISS_B200_ERR_NOMEM: return "out of memory";
    case KISS_B200_ERR_UNSUPPORTED: return "unsupported k";
    case KISS_B200_ERR_NO_DEVICE: return "no CUDA device (this library has no CPU fallback)";
    default: return "unknown error";
  }
}

const char* kiss_b200_last_error(void) { return g_last_error.c_str(); }

int kiss_b200_create(kiss_b200_ctx** out, int device) {
  if (!out) { g_last_error = "null out pointer"; return KISS_B200_ERR_ARG; }
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0 || device < 0 || device >= ndev) {
    cudaGetLastError();
    g_last_error = std::string("no usable CUDA device ") + std::to_string(device) + " (" +
                   (e != cudaSuccess ? cudaGetErrorString(e) : "device count 0 or index out of range") +
                   "); kiss_b200 has no CPU fallback";
    return KISS_B200_ERR_NO_DEVICE;
  }
  kiss_b200_ctx* c = new kiss_b200_ctx();
  c->device = device;
  try {
    CK(cudaSetDevice(device));
    // (the main stream outranks the team stream: the many small kernels of the big-group rounds are latency-bound and
    // slip in between the long-running team kernels)
    int prio_lo = 0, prio_hi = 0;
    CK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    CK(cudaStreamCreateWithPriority(&c->st, cudaStreamNonBlocking, prio_hi));
    CK(cudaStreamCreateWithFlags(&c->st_copy, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithPriority(&c->st_aux, cudaStreamNonBlocking, prio_lo));
    CK(cudaEventCreateWithFlags(&c->ev_aux, cudaEventDisableTiming));
    for (auto& ev : c->ev) CK(cudaEventCreate(&ev));
    CK(cudaHostAlloc((void**)&c->pinned, 64 * sizeof(u32) + 8 * sizeof(u64), cudaHostAllocMapped | cudaHostAllocPortable));
    cudaMemPool_t mp;
    CK(cudaDeviceGetDefaultMemPool(&mp, device));
    u64 thr = ~0ull;  // keep freed scratch cached in the pool between calls
    CK(cudaMemPoolSetAttribute(mp, cudaMemPoolAttrReleaseThreshold, &thr));
    radix_init();
    wave_init();
    // (cudaLimitMaxL2FetchGranularity 32 / 128 instead of the default measured no different on either workload)
  } catch (const CudaError& err) {
    g_last_error = err.what();
    delete c;
    return err.code;
  }
  *out = c;
  return KISS_B200_OK;
}

void kiss_b200_destroy(kiss_b200_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  free_prepared(c);
  if (c->st) cudaStreamSynchronize(c->st);
  for (auto& ev : c->ev) if (ev) cudaEventDestroy(ev);
  if (c->pinned) cudaFreeHost(c->pinned);
  if (c->st_copy) cudaStreamDestroy(c->st_copy);
  if (c->st_aux) cudaStreamDestroy(c->st_aux);
  if (c->ev_aux) cudaEventDestroy(c->ev_aux);
  if (c->st) cudaStreamDestroy(c->st);
  delete c;
}

void* kiss_b200_stream(kiss_b200_ctx* c) { return c ? (void*)c->st : nullptr; }

static int trivial_n0(u32* SA_host, kiss_b200_timings* t) {
  if (SA_host) SA_host[0] = 0;  // kiss1_core.hpp:237-238
  if (t) { memset(t, 0, sizeof(*t)); t->m = 1; }
  return KISS_B200_OK;
}

int kiss_b200_suffix_sort_dna_u32(kiss_b200_ctx* c, const uint8_t* S_host, uint64_t n, uint32_t k, int algo,
                                  uint32_t* SA_host, kiss_b200_timings* t) {
  if (!SA_host || (n && !S_host) || (algo != 0 && algo != 1)) { g_last_error = "bad argument"; return KISS_B200_ERR_ARG; }
  if (int r = check_n(n)) return r;
  if (n == 0) return c ? trivial_n0(SA_host, t) : KISS_B200_ERR_ARG;
  return guarded(c, [&] { run_host(c, S_host, n, k, algo, SA_host, n + 1, STOP_FULL, IN_CODES, t, nullptr); });
}

int kiss_b200_suffix_sort_ascii_u32(kiss_b200_ctx* c, const char* seq_host, uint64_t n, uint32_t k, int algo,
                                    uint32_t* SA_host, kiss_b200_timings* t) {
  if (!SA_host || (n && !seq_host) || (algo != 0 && algo != 1)) { g_last_error = "bad argument"; return KISS_B200_ERR_ARG; }
  if (int r = check_n(n)) return r;
  if (n == 0) return c ? trivial_n0(SA_host, t) : KISS_B200_ERR_ARG;
  return guarded(c, [&] { run_host(c, (const u8*)seq_host, n, k, algo, SA_host, n + 1, STOP_FULL, IN_ASCII, t, nullptr); });
}

int kiss_b200_suffix_sort_file_u32(kiss_b200_ctx* c, const uint8_t* file_host, uint64_t nbytes, uint32_t k, int algo,
                                   uint32_t* SA_host, uint64_t SA_capacity, uint64_t* n_out, kiss_b200_timings* t) {
  if (!c || !SA_host || !SA_capacity || !n_out || (nbytes && !file_host) || (algo != 0 && algo != 1)) {
    g_last_error = "bad argument";
    return KISS_B200_ERR_ARG;
  }
  *n_out = 0;
  if (nbytes == 0) return trivial_n0(SA_host, t);
  return guarded(c, [&] { run_host(c, file_host, nbytes, k, algo, SA_host, SA_capacity, STOP_FULL, IN_FILE, t, nullptr, n_out); });
}

int kiss_b200_ingest_dev(kiss_b200_ctx* c, const uint8_t* file_dev, uint64_t nbytes, int fasta, uint8_t* codes_dev,
                         uint64_t* n_out) {
  if (!c || !n_out || (nbytes && (!file_dev || !codes_dev))) { g_last_error = "bad argument"; return KISS_B200_ERR_ARG; }
  *n_out = 0;
  if (nbytes == 0) return KISS_B200_OK;
  return guarded(c, [&] {
    Pool pool(c->st, &c->arena);
    *n_out = do_ingest(c, pool, file_dev, nbytes, fasta != 0, codes_dev);
  });
}

int kiss_b200_suffix_sort_dna_u32_dev(kiss_b200_ctx* c, const uint8_t* S_dev, uint64_t n, uint32_t k, int algo,
                                      uint32_t* SA_dev, kiss_b200_timings* t) {
  if (!SA_dev || (n && !S_dev) || (algo != 0 && algo != 1)) { g_last_error = "bad argument"; return KISS_B200_ERR_ARG; }
  if ((((uintptr_t)S_dev | (uintptr_t)SA_dev) & 15) != 0) { g_last_error = "S_dev and SA_dev must be 16-byte aligned"; return KISS_B200_ERR_ARG; }
  if (int r = check_n(n)) return r;
  return guarded(c, [&] {
    if (n == 0) {
      CK(cudaMemsetAsync(SA_dev, 0, sizeof(u32), c->st));
      CK(cudaStreamSynchronize(c->st));
      if (t) { memset(t, 0, sizeof(*t)); t->m = 1; }
      return;
    }
    run_pipeline(c, S_dev, n, k, algo, SA_dev, STOP_FULL, t);
    free_prepared(c);
  });
}

int kiss_b200_get_lms(kiss_b200_ctx* c, const uint8_t* S_host, uint64_t n, uint32_t* lms_host, uint64_t* m_out,
                      uint64_t* count4, uint64_t* lms_count4) {
  if (n && !S_host) { g_last_error = "bad argument"; return KISS_B200_ERR_ARG; }
  if (int r = check_n(n)) return r;
  if (n == 0) {
    if (lms_host) lms_host[0] = 0;
    if (m_out) *m_out = 1;
    for (int s = 0; s < 4; s++) { if (count4) count4[s] = 0; if (lms_count4) lms_count4[s] = 0; }
    return KISS_B200_OK;
  }
  return guarded(c, [&] {
    run_host(c, S_host, n, 0xFFFFFFFFu, 0, lms_host, 0, STOP_LMS, false, nullptr, m_out);
    for (int s = 0; s < 4; s++) {
      if (count4) count4[s] = c->prep.counters[s];
      if (lms_count4) lms_count4[s] = c->prep.counters[4 + s];
    }
    free_prepared(c);
  });
}

int kiss_b200_sorted_lms(kiss_b200_ctx* c, const uint8_t* S_host, uint64_t n, uint32_t k, int algo,
                         uint32_t* sorted_host, uint64_t* m_out) {
  if (!sorted_host || (n && !S_host) || (algo != 0 && algo != 1)) { g_last_error = "bad argument"; return KISS_B200_ERR_ARG; }
  if (int r = check_n(n)) return r;
  if (n == 0) { sorted_host[0] = 0; if (m_out) *m_out = 1; return KISS_B200_OK; }
  return guarded(c, [&] {
    run_host(c, S_host, n, k, algo, sorted_host, 0, STOP_SORT, false, nullptr, m_out);
    free_prepared(c);
  });
}

int kiss_b200_put_lms(kiss_b200_ctx* c, const uint8_t* S_host, uint64_t n, uint32_t k, int algo, uint32_t* SA_host) {
  if (!SA_host || (n && !S_host) || (algo != 0 && algo != 1)) { g_last_error = "bad argument"; return KISS_B200_ERR_ARG; }
  if (int r = check_n(n)) return r;
  if (n == 0) { SA_host[0] = 0; return KISS_B200_OK; }
  return guarded(c, [&] {
    run_host(c, S_host, n, k, algo, SA_host, 0, STOP_PUT, false, nullptr, nullptr);
    free_prepared(c);
  });
}

int kiss_b200_induce_from_sorted_lms(kiss_b200_ctx* c, const uint8_t* S_host, uint64_t n, const uint32_t* sorted_host,
                                     uint64_t m, uint32_t* SA_host) {
  if (!SA_host || !sorted_host || (n && !S_host) || m == 0) { g_last_error = "bad argument"; return KISS_B200_ERR_ARG; }
  if (int r = check_n(n)) return r;
  if (n == 0) { SA_host[0] = 0; return KISS_B200_OK; }
  return guarded(c, [&] {
    cudaStream_t st = c->st;
    begin_call(c);
    Pool pool(st, &c->arena);
    u8* dS = pool.alloc<u8>(n + 16);
    CK(cudaMemcpyAsync(dS, S_host, n, cudaMemcpyHostToDevice, st));
    do_prepare(c, dS, n, 0xFFFFFFFFu, 0, nullptr, nullptr);
    if (c->prep.m1 + 1 != m) throw CudaError(KISS_B200_ERR_ARG, "sorted list length differs from the LMS count of the text");
    u32* dsorted = pool.alloc<u32>(m);
    CK(cudaMemcpyAsync(dsorted, sorted_host, m * sizeof(u32), cudaMemcpyHostToDevice, st));
    u32* dSA = pool.alloc<u32>(n + 1);
    Buckets bk = make_buckets(c->prep.counters, n);
    u32* dCTX = pool.alloc<u32>(n + 1);
    InduceStats is;
    do_induce(c, pool, c->prep.T, n, bk, dsorted + 1, dSA, dCTX, &is);
    CK(cudaMemcpyAsync(SA_host, dSA, (n + 1) * sizeof(u32), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    free_prepared(c);
  });
}

int kiss_b200_prepare_dev(kiss_b200_ctx* c, const uint8_t* S_dev, uint64_t n, uint32_t k, uint64_t* m_out) {
  if (!S_dev || n == 0) { g_last_error = "bad argument"; return KISS_B200_ERR_ARG; }
  if ((uintptr_t)S_dev & 15) { g_last_error = "S_dev must be 16-byte aligned"; return KISS_B200_ERR_ARG; }
  if (int r = check_n(n)) return r;
  return guarded(c, [&] {
    do_prepare(c, S_dev, n, k, 0, nullptr, nullptr);
    if (m_out) *m_out = c->prep.m1 + 1;
  });
}

int kiss_b200_lms_keys_dev(kiss_b200_ctx* c, const uint32_t** keys_dev, const uint32_t** pos_dev, uint64_t* n_long) {
  if (!c || !c->prep.valid) { g_last_error = "no prepared text"; return KISS_B200_ERR_ARG; }
  if (keys_dev) *keys_dev = c->prep.key;
  if (pos_dev) *pos_dev = c->prep.pos;
  if (n_long) *n_long = c->prep.n_long;
  return KISS_B200_OK;
}

int kiss_b200_partition_lms_dev(kiss_b200_ctx* c, const uint32_t* splitters_host, int n_split, uint64_t lo,
                                uint64_t hi, uint32_t* out_pos_dev, uint64_t* counts_host) {
  if (!c || !c->prep.valid) { g_last_error = "no prepared text"; return KISS_B200_ERR_ARG; }
  if (n_split < 0 || n_split > 63 || lo > hi || hi > c->prep.n_long || (hi > lo && !out_pos_dev) || !counts_host ||
      (n_split && !splitters_host)) {
    g_last_error = "bad argument";
    return KISS_B200_ERR_ARG;
  }
  return guarded(c, [&] {
    cudaStream_t st = c->st;
    Pool pool(st, &c->arena);
    u32 N = (u32)(hi - lo);
    u32* spl = pool.alloc<u32>(64);
    u32* counts = pool.alloc<u32>(65);
    CK(cudaMemsetAsync(counts, 0, 65 * sizeof(u32), st));
    if (n_split) CK(cudaMemcpyAsync(spl, splitters_host, n_split * sizeof(u32), cudaMemcpyHostToDevice, st));
    u32* dest = pool.alloc<u32>(N);
    u32* dest_alt = pool.alloc<u32>(N);
    u32* vals = pool.alloc<u32>(N);
    u32* vals_alt = pool.alloc<u32>(N);
    if (N) CK(cudaMemcpyAsync(vals, c->prep.pos + lo, N * sizeof(u32), cudaMemcpyDeviceToDevice, st));
    dest_by_splitters(c->prep.key + lo, N, spl, (u32)n_split, dest, counts, st);
    u64 nt = radix_tiles(N);
    u32* H = pool.alloc<u32>(256 * nt);
    u32* partials = pool.alloc<u32>(scan_partials_len(256 * nt));
    radix_sort_u32(dest, vals, dest_alt, vals_alt, N, 0, 8, H, partials, st);  // one stable pass on the destination
    if (N) CK(cudaMemcpyAsync(out_pos_dev, vals, N * sizeof(u32), cudaMemcpyDeviceToDevice, st));
    u32 hc[65];
    CK(cudaMemcpyAsync(hc, counts, 65 * sizeof(u32), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    for (int d = 0; d <= n_split; d++) counts_host[d] = hc[d];
  });
}

int kiss_b200_sort_lms_subset_dev(kiss_b200_ctx* c, uint32_t* pos_dev, uint64_t count, kiss_b200_timings* t) {
  if (!c || !c->prep.valid) { g_last_error = "no prepared text"; return KISS_B200_ERR_ARG; }
  if (count && !pos_dev) { g_last_error = "bad argument"; return KISS_B200_ERR_ARG; }
  if (count > 0x80000000ull) { g_last_error = "subset too large"; return KISS_B200_ERR_TOO_LARGE; }
  return guarded(c, [&] {
    cudaStream_t st = c->st;
    Pool pool(st, &c->arena);
    u64 l0 = g_launches;
    CK(cudaEventRecord(c->ev[3], st));
    u32* key = pool.alloc<u32>(count);
    gather_keys(c->prep.sp.T, pos_dev, (u32)count, key, st);
    u32* sorted;
    SortStats ss;
    sort_long_lms(c, pool, key, pos_dev, (u32)count, c->prep.sp, &sorted, &ss);
    if (sorted != pos_dev && count)
      CK(cudaMemcpyAsync(pos_dev, sorted, count * sizeof(u32), cudaMemcpyDeviceToDevice, st));
    CK(cudaEventRecord(c->ev[4], st));
    CK(cudaEventSynchronize(c->ev[4]));
    if (t) {
      memset(t, 0, sizeof(*t));
      t->lms_sort = elapsed_s(c, 3, 4);
      t->total_device = t->lms_sort;
      t->n = c->prep.n; t->m = c->prep.m1 + 1;
      t->lms_sort_tied = ss.tied; t->lms_sort_rounds = ss.rounds;
      t->kernel_launches = g_launches - l0;
    }
  });
}

static int finish_from_sorted(kiss_b200_ctx* c, const uint32_t* sorted_long_dev, const uint32_t* ctx_long_dev,
                              uint64_t n_long, uint32_t* SA_dev, kiss_b200_timings* t) {
  if (!c || !c->prep.valid) { g_last_error = "no prepared text"; return KISS_B200_ERR_ARG; }
  if (!SA_dev || (n_long && !sorted_long_dev) || n_long != c->prep.n_long || ((uintptr_t)SA_dev & 15)) { g_last_error = "bad argument"; return KISS_B200_ERR_ARG; }
  return guarded(c, [&] {
    cudaStream_t st = c->st;
    Pool pool(st, &c->arena);
    auto& P = c->prep;
    u64 l0 = g_launches;
    u32 ns = (u32)(P.m1 - P.n_long);
    CK(cudaEventRecord(c->ev[4], st));
    SortedLms sl;
    sl.longs = const_cast<u32*>(sorted_long_dev);
    sl.nl = P.n_long;
    do_shorts(c, pool, P.pos + P.n_long, ns, sorted_long_dev, P.n_long, P.sp, &sl);
    Buckets bk = make_buckets(P.counters, P.n);
    u32* dCTX = pool.alloc<u32>(P.n + 1);
    // the LMS waves read (long suffixes, short ones inserted on the fly): neither the positions nor the contexts
    // fetched by the ranks that sorted the ranges are merged into new arrays
    const u32* lms = sl.longs;
    const u32* lms_ctx = ctx_long_dev;
    CK(cudaEventRecord(c->ev[5], st));
    InduceStats is;
    do_induce(c, pool, P.T, P.n, bk, lms, SA_dev, dCTX, &is, nullptr, lms_ctx, sl.shorts, sl.ins, sl.ns);
    CK(cudaEventRecord(c->ev[6], st));
    CK(cudaEventSynchronize(c->ev[6]));
    if (t) {
      memset(t, 0, sizeof(*t));
      t->put_lms = elapsed_s(c, 4, 5);
      t->induce = elapsed_s(c, 5, 6);
      t->total_device = t->put_lms + t->induce;
      t->n = P.n; t->m = P.m1 + 1;
      t->waves_L = is.waves_L; t->waves_S = is.waves_S;
      t->kernel_launches = g_launches - l0;
    }
  });
}

int kiss_b200_finish_from_sorted_dev(kiss_b200_ctx* c, const uint32_t* sorted_long_dev, uint64_t n_long,
                                     uint32_t* SA_dev, kiss_b200_timings* t) {
  return finish_from_sorted(c, sorted_long_dev, nullptr, n_long, SA_dev, t);
}

int kiss_b200_finish_from_sorted_ctx_dev(kiss_b200_ctx* c, const uint32_t* sorted_long_dev, const uint32_t* ctx_long_dev,
                                         uint64_t n_long, uint32_t* SA_dev, kiss_b200_timings* t) {
  return finish_from_sorted(c, sorted_long_dev, ctx_long_dev, n_long, SA_dev, t);
}

int kiss_b200_lms_contexts_dev(kiss_b200_ctx* c, const uint32_t* pos_dev, uint64_t count, uint32_t* ctx_out_dev) {
  if (!c || !c->prep.valid) { g_last_error = "no prepared text"; return KISS_B200_ERR_ARG; }
  if (count && (!pos_dev || !ctx_out_dev)) { g_last_error = "bad argument"; return KISS_B200_ERR_ARG; }
  return guarded(c, [&] {
    lms_contexts(c->prep.T, pos_dev, count, ctx_out_dev, c->st);  // (guarded() waits for the stream before returning)
  });
}

int kiss_b200_profile_enable(kiss_b200_ctx* c, int on) {
  if (!c) { g_last_error = "null context"; return KISS_B200_ERR_ARG; }
  c->prof.on = on != 0;
  c->prof.reset();
  return KISS_B200_OK;
}

int kiss_b200_profile_read(kiss_b200_ctx* c, int idx, const char** name, double* ms, uint64_t* launches) {
  if (!c || idx < 0 || idx >= PK_COUNT) return KISS_B200_ERR_ARG;
  if (name) *name = prof_name(idx);
  if (ms) *ms = c->prof.ms[idx];
  if (launches) *launches = c->prof.launches[idx];
  return KISS_B200_OK;
}

int kiss_b200_profile_alg_bytes(kiss_b200_ctx* c, int idx, uint64_t* bytes) {
  if (!c || !bytes || idx < 0 || idx >= PK_COUNT) return KISS_B200_ERR_ARG;
  *bytes = c->prof.bytes[idx];
  return KISS_B200_OK;
}

uint64_t kiss_b200_fmindex_image_bytes(uint64_t n, uint32_t sa_intv) { return sa_intv ? fm_layout(n, sa_intv).total : 0; }

static int fm_args_ok(kiss_b200_ctx* c, const void* in, uint64_t n, int algo, uint32_t sa_intv, const void* image,
                      const uint64_t* image_bytes) {
  if (!c || !image || !image_bytes || (n && !in) || (algo != 0 && algo != 1) || sa_intv == 0) {
    g_last_error = "bad argument";
    return KISS_B200_ERR_ARG;
  }
  return 0;
}

int kiss_b200_fmindex_build_dna_u32(kiss_b200_ctx* c, const uint8_t* S_host, uint64_t n, uint32_t k, int algo,
                                    uint32_t sa_intv, uint8_t* image_host, uint64_t capacity, uint64_t* image_bytes,
                                    kiss_b200_timings* t, double* fm_seconds) {
  if (int r = fm_args_ok(c, S_host, n, algo, sa_intv, image_host, image_bytes)) return r;
  if (int r = check_n(n)) return r;
  return guarded(c, [&] {
    run_fm_host(c, S_host, n, IN_CODES, k, algo, sa_intv, image_host, capacity, image_bytes, nullptr, t, fm_seconds);
  });
}

int kiss_b200_fmindex_build_file_u32(kiss_b200_ctx* c, const uint8_t* file_host, uint64_t nbytes, uint32_t k, int algo,
                                     uint32_t sa_intv, uint8_t* image_host, uint64_t capacity, uint64_t* image_bytes,
                                     uint64_t* n_out, kiss_b200_timings* t, double* fm_seconds) {
  if (int r = fm_args_ok(c, file_host, nbytes, algo, sa_intv, image_host, image_bytes)) return r;
  return guarded(c, [&] {
    run_fm_host(c, file_host, nbytes, IN_FILE, k, algo, sa_intv, image_host, capacity, image_bytes, n_out, t, fm_seconds);
  });
}

int kiss_b200_fmindex_from_sa_dev(kiss_b200_ctx* c, const uint8_t* S_dev, uint64_t n, const uint32_t* SA_dev,
                                  uint32_t sa_intv, uint8_t* image_host, uint64_t capacity, uint64_t* image_bytes,
                                  double* fm_seconds) {
  if (int r = fm_args_ok(c, S_dev, n, 0, sa_intv, image_host, image_bytes)) return r;
  if (!SA_dev || ((uintptr_t)S_dev & 15)) { g_last_error = "bad argument (SA_dev null or S_dev not 16-byte aligned)"; return KISS_B200_ERR_ARG; }
  if (int r = check_n(n)) return r;
  return guarded(c, [&] {
    const FmLayout L = fm_layout(n, sa_intv);
    *image_bytes = L.total;
    if (L.total > capacity) throw CudaError(KISS_B200_ERR_ARG, "image buffer smaller than kiss_b200_fmindex_image_bytes(n, sa_intv)");
    Pool pool(c->st, &c->arena);
    u64* T = pool.alloc<u64>(packed_words(n));
    if (n) pack_codes(S_dev, n, T, c->st);
    float ms = 0;
    do_fmindex(c, pool, T, n, SA_dev, nullptr, sa_intv, image_host, &ms);
    if (fm_seconds) *fm_seconds = ms * 1e-3;
  });
}

uint64_t kiss_b200_launch_count(void) { return g_launches; }

int kiss_b200_sa_checksum_dev(kiss_b200_ctx* c, const uint32_t* SA_dev, uint64_t count, uint64_t* out) {
  if (!out || (count && !SA_dev)) { g_last_error = "bad argument"; return KISS_B200_ERR_ARG; }
  return guarded(c, [&] {
    Pool pool(c->st, &c->arena);
    u64* d = pool.alloc<u64>(1);
    mixed_sum(SA_dev, count, d, c->st);
    CK(cudaMemcpyAsync(out, d, sizeof(u64), cudaMemcpyDeviceToHost, c->st));
    CK(cudaStreamSynchronize(c->st));
  });
}

int kiss_b200_verify_sa_dev(kiss_b200_ctx* c, const uint8_t* S_dev, uint64_t n, const uint32_t* SA_dev, uint32_t k,
                            uint64_t counts[4]) {
  if (!counts || !SA_dev || (n && !S_dev) || ((uintptr_t)S_dev & 15)) { g_last_error = "bad argument (null pointer or S_dev not 16-byte aligned)"; return KISS_B200_ERR_ARG; }
  if (int rc = check_n(n)) return rc;
  return guarded(c, [&] {
    Pool pool(c->st, &c->arena);
    u64* T = pool.alloc<u64>(packed_words(n));
    if (n) pack_codes(S_dev, n, T, c->st);
    else CK(cudaMemsetAsync(T, 0, packed_words(0) * sizeof(u64), c->st));
    u32* bitmap = pool.alloc<u32>((n + 1 + 31) / 32);
    u64* d = pool.alloc<u64>(4);
    verify_sa(T, n, SA_dev, n + 1, k, 4096u, bitmap, d, c->st);
    CK(cudaMemcpyAsync(counts, d, 4 * sizeof(u64), cudaMemcpyDeviceToHost, c->st));
    CK(cudaStreamSynchronize(c->st));
  });
}

int kiss_b200_verify_sa_bytes_dev(kiss_b200_ctx* c, const uint8_t* S_dev, uint64_t n, const uint32_t* SA_dev, uint32_t k,
                                  uint64_t counts[4]) {
  if (!counts || !SA_dev || (n && !S_dev)) { g_last_error = "bad argument"; return KISS_B200_ERR_ARG; }
  if (int rc = check_n(n)) return rc;
  return guarded(c, [&] {
    Pool pool(c->st, &c->arena);
    u32* bitmap = pool.alloc<u32>((n + 1 + 31) / 32);
    u64* d = pool.alloc<u64>(4);
    verify_sa_bytes(S_dev, n, SA_dev, n + 1, k, 131072u, bitmap, d, c->st);
    CK(cudaMemcpyAsync(counts, d, 4 * sizeof(u64), cudaMemcpyDeviceToHost, c->st));
    CK(cudaStreamSynchronize(c->st));
  });
}

int kiss_b200_synth_uniform_dev(kiss_b200_ctx* c, uint8_t* S_dev, uint64_t n, uint64_t seed) {
  if (!S_dev && n) { g_last_error = "bad argument"; return KISS_B200_ERR_ARG; }
  return guarded(c, [&] {
    synth_uniform(S_dev, n, seed, c->st);
    CK(cudaStreamSynchronize(c->st));
  });
}

}  // extern "C"
