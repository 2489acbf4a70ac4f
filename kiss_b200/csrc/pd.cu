// pd.cu — K3': the PREFIX_DOUBLING LMS sort (kiss2_core.hpp:728-798) as rank doubling on the device.
//
// The reference chops every LMS substring into words of 16 bases (encode_S, kiss2_core.hpp:20-100: a word starts at the
// LMS position and every 16 bases after it while it is before the next LMS position; it reads 16 bases whatever follows,
// zero past the end), orders the words by value with a stable LSD radix sort in which the words that overlap the text
// end come first in reverse index order and are groups of their own (initialize_sa / initialize_rank, :113-249), and
// then doubles: groups are ordered by rank[index + offset] for offset = 1, 2, 4, ... < k over WORD indices, split where
// that rank changes, and singletons leave the work list (:764-785).  The order of the LMS-start words is the LMS order
// (place_back_lms, :624-726).
//
// Here: the same words, the same initial order (radix.cu's stable LSD sort on an input permuted like :130-133), and per
// round one composite key [group ordinal | rank at index + offset] per ACTIVE element, sorted with the stable radix sort;
// new group heads where the composite changes; ranks = head slot + 1 as in the reference (a rank never changes once its
// group is a singleton); singletons are compacted away.  O(log2 of the longest repeat in words) rounds whatever the
// text looks like — the pivot rounds of lms_sort.cu cost O(LCP / 32) word loads per element, which has no bound on texts
// with Mbp-scale exact duplicates.  With a bounded k the loop stops at offset >= k like the reference's, and what is still
// tied keeps ascending word index = ascending text position (every step is stable): exactly the classes of
// tests/kiss2_words.py, in canonical order — where the reference leaves them in an order that depends on its thread
// count.
#include "common.cuh"
#include "kernels.h"
#include "team.cuh"

namespace kiss {

#define PD_T 256

// cnt[i] = words of LMS substring i (kiss2_core.hpp:45-47)
__global__ void k_pd_count(const u32* __restrict__ lms, u32 m1, u64 n, u32* __restrict__ cnt) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m1) return;
  const u64 next = i + 1 < m1 ? lms[i + 1] : n;
  cnt[i] = (u32)((next - lms[i] + 15) >> 4);
}

// word starts, LMS-start flags and the first 16 bases of every word, written at the word's INPUT SLOT of the initial
// sort: the `ntail` words that overlap the text end (the last ones by index) first, in reverse index order (:130-133)
__global__ void k_pd_words(const u32* __restrict__ lms, const u32* __restrict__ first, const u32* __restrict__ cnt, u32 m1,
                           const u64* __restrict__ T, u32 W, u32 ntail, u32* __restrict__ wstart, u32* __restrict__ isstart /* bit per word, zeroed */,
                           u32* __restrict__ key_in, u32* __restrict__ idx_in) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m1) return;
  const u32 f = first[i], c = cnt[i];
  const u64 p0 = lms[i];
  for (u32 j = 0; j < c; j++) {
    const u32 w = f + j;
    const u64 p = p0 + 16ull * j;
    wstart[w] = (u32)p;
    if (j == 0) atomicOr(&isstart[w >> 5], 1u << (w & 31u));  // (a bitmap: W / 8 bytes stay in L2 for the final gathers)
    const u32 slot = w >= W - ntail ? W - 1 - w : w + ntail;
    key_in[slot] = (u32)(word_at(T, p) >> 32);
    idx_in[slot] = w;
  }
}

// number of words starting after n - 16 (:81-84); wstart is ascending
__global__ void k_pd_ntail(const u32* __restrict__ lms, const u32* __restrict__ cnt, u32 m1, u64 n, u32* __restrict__ out) {
  // only the last few LMS substrings can reach the last 16 bases: walk back from the end
  if (threadIdx.x || blockIdx.x) return;
  u32 t = 0;
  for (u32 i = m1; i-- > 0;) {
    const u64 p0 = lms[i];
    const u32 c = cnt[i];
    const u64 last = p0 + 16ull * (c - 1);
    if (last + 16 <= n) break;  // every word of this and of the earlier substrings fits
    for (u32 j = 0; j < c; j++) t += (p0 + 16ull * j + 16 > n) ? 1u : 0u;
  }
  *out = t;
}

// initial heads (:215-217): the value changes, or the element before is a tail word
__global__ void k_pd_heads0(const u32* __restrict__ key, const u32* __restrict__ idx, u32 W, u32 ntail, u32* __restrict__ head) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= W) return;
  head[i] = (i == 0 || key[i] != key[i - 1] || idx[i - 1] >= W - ntail) ? 1u : 0u;
}

// heads after a round: the composite key changes (it contains the group ordinal, so old boundaries stay)
__global__ void k_pd_heads(const u64* __restrict__ comp, u32 A, u32* __restrict__ head) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A) return;
  head[i] = (i == 0 || comp[i] != comp[i - 1]) ? 1u : 0u;
}

// ord = exclusive scan of head: the head of ordinal o sits at slot headslot[o]
__global__ void k_pd_headslot(const u32* __restrict__ head, const u32* __restrict__ ord, u32 A, u32* __restrict__ headslot) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < A && head[i]) headslot[ord[i]] = i;
}

// rank of every element = global position of its group's head + 1 (:226-248, :378-462); gslot == nullptr: slot = position
// `isnew` (nullable): bit per slot, set where this round made a new head.  A group whose head was a head before starts
// where its elements' old group started, so their rank is what it was: the random 4-byte write is skipped.
__global__ void k_pd_rank(const u32* __restrict__ elem, const u32* __restrict__ head, const u32* __restrict__ ord,
                          const u32* __restrict__ headslot, const u32* __restrict__ gslot, const u32* __restrict__ isnew, u32 A,
                          u32* __restrict__ rank) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A) return;
  const u32 hs = headslot[ord[i] + head[i] - 1u];
  if (isnew && !((isnew[hs >> 5] >> (hs & 31u)) & 1u)) return;
  rank[elem[i]] = (gslot ? gslot[hs] : hs) + 1u;
}

// keep[i] = the element shares its group with another one
__global__ void k_pd_keep(const u32* __restrict__ head, u32 A, u32* __restrict__ keep) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A) return;
  const bool single = head[i] && (i + 1 == A || head[i + 1]);
  keep[i] = single ? 0u : 1u;
}

// singletons are final: sa_final[global position] = word; the others move to the compacted work list
__global__ void k_pd_compact(const u32* __restrict__ elem, const u32* __restrict__ gslot, const u32* __restrict__ head,
                             const u32* __restrict__ keep, const u32* __restrict__ dst, u32 A, u32* __restrict__ sa_final,
                             u32* __restrict__ elem2, u32* __restrict__ gslot2, u32* __restrict__ head2) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A) return;
  const u32 g = gslot ? gslot[i] : i;
  if (!keep[i]) {
    sa_final[g] = elem[i];
  } else {
    const u32 d = dst[i];
    elem2[d] = elem[i];
    gslot2[d] = g;
    head2[d] = head[i];
  }
}

// composite key of a round: [group ordinal | rank[index + offset]] (rank 0 past the last word, :102-111)
__global__ void k_pd_key2(const u32* __restrict__ elem, const u32* __restrict__ head, const u32* __restrict__ ord,
                          const u32* __restrict__ rank, u32 A, u32 W, u32 off, int rb, u64* __restrict__ comp) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A) return;
  const u64 w = (u64)elem[i] + off;
  const u32 r = w < W ? rank[w] : 0u;
  comp[i] = ((u64)(ord[i] + head[i] - 1u) << rb) | r;  // rb = bits of a rank: the sort runs over rb + bits(groups) bits
}

// what is still tied when the loop ends keeps its (stable) order
__global__ void k_pd_flush(const u32* __restrict__ elem, const u32* __restrict__ gslot, u32 A, u32* __restrict__ sa_final) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < A) sa_final[gslot[i]] = elem[i];
}

// LMS-start words in final order (:690-725): flag = the word starts an LMS substring, val = its text position (> 0) or 0;
// after the scan of the flags the positions go out in order
__global__ void k_pd_startflags(const u32* __restrict__ sa_final, const u32* __restrict__ isstart, const u32* __restrict__ wstart,
                                u32 W, u32* __restrict__ flag, u32* __restrict__ val) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= W) return;
  const u32 w = sa_final[i];
  const u32 s = (isstart[w >> 5] >> (w & 31u)) & 1u;
  flag[i] = s;
  val[i] = s ? wstart[w] : 0u;
}
__global__ void k_pd_emit(const u32* __restrict__ val, const u32* __restrict__ dst, u32 W, u32* __restrict__ out) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= W) return;
  const u32 v = val[i];
  if (v) out[dst[i]] = v;
}

// ---- segmented sort of a round.  The active elements are already grouped (a group = a run of slots), so a round only
// has to order every group by its second rank key2 = rank[index + offset]: groups of at most 8 by one thread, up to 256
// by a warp team, up to 2048 by a block team (the register-resident bitonic network of team.cuh on [key2 | slot], stable
// through the slot field), all in ONE pass over the data — the global radix sort of the composite key [group | key2]
// needs 7 passes (512 Mbp of 512 near-identical copies, -k -1: 27 ms per round against ~5).  A round with a larger
// group takes the radix sort (the counters tell the host).
__global__ void k_pd_key2_plain(const u32* __restrict__ elem, const u32* __restrict__ rank, u32 A, u32 W, u32 off,
                                u32* __restrict__ key2) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A) return;
  const u64 w = (u64)elem[i] + off;
  key2[i] = w < W ? rank[w] : 0u;
}

// one thread per group: orders the groups of <= 8 in place, lists the others by size class (7 lists of list_cap entries);
// counters: [0..6] groups per class, [7] groups > 2048
__global__ void k_pd_seg_small(const u32* __restrict__ headslot, u32 ngroups, u32 A, u32* __restrict__ elem,
                               u32* __restrict__ key2, u32* __restrict__ lists, u32 list_cap, u32* __restrict__ counters) {
  const u32 g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= ngroups) return;
  const u32 start = headslot[g], end = g + 1 < ngroups ? headslot[g + 1] : A, size = end - start;
  if (size <= 8) {
    u32 k[8], e[8];
#pragma unroll
    for (u32 a = 0; a < 8; a++) {
      k[a] = a < size ? key2[start + a] : 0xFFFFFFFFu;
      e[a] = a < size ? elem[start + a] : 0u;
    }
    // stable insertion network on (key, input order): a compare-exchange only swaps on strictly greater
#pragma unroll
    for (u32 a = 1; a < 8; a++)
#pragma unroll
      for (u32 b = a; b > 0; b--) {
        const bool sw = b < size && k[b - 1] > k[b];
        const u32 tk = k[b], te = e[b];
        k[b] = sw ? k[b - 1] : k[b];   e[b] = sw ? e[b - 1] : e[b];
        k[b - 1] = sw ? tk : k[b - 1]; e[b - 1] = sw ? te : e[b - 1];
      }
#pragma unroll
    for (u32 a = 0; a < 8; a++)
      if (a < size) { key2[start + a] = k[a]; elem[start + a] = e[a]; }
  } else if (size <= 2048) {
    // size classes of the teams: <= 32, 64, 128, 256 (a warp with 1..8 keys per lane), <= 512, 1024, 2048 (a block)
    const u32 cls = size <= 32 ? 0u : size <= 64 ? 1u : size <= 128 ? 2u : size <= 256 ? 3u : size <= 512 ? 4u : size <= 1024 ? 5u : 6u;
    lists[(u64)cls * list_cap + atomicAdd(&counters[cls], 1u)] = g;
  } else {
    // larger than a block team: listed after the seven class lists (at most A / 2049 of them) for the radix sort of the
    // subset (pd_big_*)
    lists[7ull * list_cap + atomicAdd(&counters[7], 1u)] = g;
  }
}

// ---- groups larger than a block team: their elements are copied out with the composite key [group ordinal | key2],
// radix-sorted (stable), and copied back.  bsize = sizes of the listed groups (to be scanned into boff).
__global__ void k_pd_big_sizes(const u32* __restrict__ biglist, u32 nb, const u32* __restrict__ headslot, u32 ngroups, u32 A,
                               u32* __restrict__ bsize) {
  const u32 j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nb) return;
  const u32 g = biglist[j];
  bsize[j] = (g + 1 < ngroups ? headslot[g + 1] : A) - headslot[g];
}
__global__ void k_pd_big_fill(const u32* __restrict__ biglist, const u32* __restrict__ boff, u32 nb, u32 B,
                              const u32* __restrict__ headslot, const u32* __restrict__ elem, const u32* __restrict__ key2,
                              int rb, u64* __restrict__ comp, u32* __restrict__ val, u32* __restrict__ bslot) {
  const u32 t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= B) return;
  u32 lo = 0, hi = nb;  // last listed group with boff <= t
  while (hi - lo > 1) {
    const u32 mid = lo + (hi - lo) / 2;
    if (boff[mid] <= t) lo = mid; else hi = mid;
  }
  const u32 slot = headslot[biglist[lo]] + (t - boff[lo]);
  comp[t] = ((u64)lo << rb) | key2[slot];
  val[t] = elem[slot];
  bslot[t] = slot;  // (a function of t alone: position t of the sorted subset goes back to the same slot)
}
__global__ void k_pd_big_writeback(const u64* __restrict__ comp, const u32* __restrict__ val, const u32* __restrict__ bslot,
                                   u32 B, int rb, u32* __restrict__ elem, u32* __restrict__ key2) {
  const u32 t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= B) return;
  const u32 slot = bslot[t];
  elem[slot] = val[t];
  key2[slot] = (u32)(comp[t] & ((1ull << rb) - 1ull));
}

template <int NW, int R>
__global__ void __launch_bounds__(NW == 1 ? 256 : NW * 32) k_pd_seg_team(const u32* __restrict__ list, u32 nlist,
                                                                         const u32* __restrict__ headslot, u32 ngroups, u32 A,
                                                                         u32* __restrict__ elem, u32* __restrict__ key2) {
  constexpr u32 P = NW * R * 32, TEAMS = NW == 1 ? 8 : 1;
  __shared__ u64 s_x[NW > 1 ? P : 1];
  __shared__ u32 s_e[TEAMS * P];
  const u32 lane = threadIdx.x & 31u, wb = threadIdx.x >> 5;
  const u32 team = NW == 1 ? wb : 0u, w = NW == 1 ? 0u : wb;
  u32* se = s_e + team * P;
  for (u32 gi = blockIdx.x * TEAMS + team; gi < nlist; gi += gridDim.x * TEAMS) {
    const u32 g = list[gi];
    const u32 start = headslot[g], end = g + 1 < ngroups ? headslot[g + 1] : A, size = end - start;
    u64 x[R], xl[R];
#pragma unroll
    for (int r = 0; r < R; r++) {
      const u32 i = ((w * R + r) << 5) | lane;
      x[r] = i < size ? (((u64)key2[start + i] << MK_SLOT_BITS) | i) : ~0ull;
      xl[r] = 0;
      se[i] = i < size ? elem[start + i] : 0u;
    }
    team_sync<NW>();
    team_sort<NW, R, false>(x, xl, w, lane, s_x, s_x);
#pragma unroll
    for (int r = 0; r < R; r++) {
      const u32 i = ((w * R + r) << 5) | lane;
      if (i < size) {
        key2[start + i] = (u32)(x[r] >> MK_SLOT_BITS);
        elem[start + i] = se[(u32)x[r] & 0x7FFu];
      }
    }
    team_sync<NW>();
  }
}

// heads after a segmented round: the old group starts, and where key2 changes
__global__ void k_pd_heads_seg(const u32* __restrict__ key2, u32 A, u32* __restrict__ head, u32* __restrict__ isnew) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31u;
  bool nh = false;
  if (i < A && i > 0 && !head[i] && key2[i] != key2[i - 1]) {
    head[i] = 1u;
    nh = true;
  }
  const u32 m = __ballot_sync(0xffffffffu, nh);  // (32 consecutive slots per warp: one bitmap word)
  if (lane == 0 && i < A) isnew[i >> 5] = m;
}

#define PD_GRID(n) (unsigned)(((u64)(n) + PD_T - 1) / PD_T)
#define PD_LAUNCH(kern, n, ...)                              \
  do {                                                       \
    if (n) {                                                 \
      kern<<<PD_GRID(n), PD_T, 0, st>>>(__VA_ARGS__);        \
      CK_LAUNCH();                                           \
      g_launches++;                                          \
    }                                                        \
  } while (0)

void pd_count(const u32* lms, u32 m1, u64 n, u32* cnt, cudaStream_t st) { PD_LAUNCH(k_pd_count, m1, lms, m1, n, cnt); }
void pd_ntail(const u32* lms, const u32* cnt, u32 m1, u64 n, u32* out, cudaStream_t st) {
  k_pd_ntail<<<1, 32, 0, st>>>(lms, cnt, m1, n, out);
  CK_LAUNCH(); g_launches++;
}
void pd_words(const u32* lms, const u32* first, const u32* cnt, u32 m1, const u64* T, u32 W, u32 ntail, u32* wstart,
              u32* isstart, u32* key_in, u32* idx_in, cudaStream_t st) {
  PD_LAUNCH(k_pd_words, m1, lms, first, cnt, m1, T, W, ntail, wstart, isstart, key_in, idx_in);
}
void pd_heads0(const u32* key, const u32* idx, u32 W, u32 ntail, u32* head, cudaStream_t st) {
  PD_LAUNCH(k_pd_heads0, W, key, idx, W, ntail, head);
}
void pd_heads(const u64* comp, u32 A, u32* head, cudaStream_t st) { PD_LAUNCH(k_pd_heads, A, comp, A, head); }
void pd_headslot(const u32* head, const u32* ord, u32 A, u32* headslot, cudaStream_t st) {
  PD_LAUNCH(k_pd_headslot, A, head, ord, A, headslot);
}
void pd_rank(const u32* elem, const u32* head, const u32* ord, const u32* headslot, const u32* gslot, const u32* isnew, u32 A,
             u32* rank, cudaStream_t st) {
  PD_LAUNCH(k_pd_rank, A, elem, head, ord, headslot, gslot, isnew, A, rank);
}
void pd_keep(const u32* head, u32 A, u32* keep, cudaStream_t st) { PD_LAUNCH(k_pd_keep, A, head, A, keep); }
void pd_compact(const u32* elem, const u32* gslot, const u32* head, const u32* keep, const u32* dst, u32 A, u32* sa_final,
                u32* elem2, u32* gslot2, u32* head2, cudaStream_t st) {
  PD_LAUNCH(k_pd_compact, A, elem, gslot, head, keep, dst, A, sa_final, elem2, gslot2, head2);
}
void pd_key2(const u32* elem, const u32* head, const u32* ord, const u32* rank, u32 A, u32 W, u32 off, int rb, u64* comp,
             cudaStream_t st) {
  PD_LAUNCH(k_pd_key2, A, elem, head, ord, rank, A, W, off, rb, comp);
}
void pd_flush(const u32* elem, const u32* gslot, u32 A, u32* sa_final, cudaStream_t st) {
  PD_LAUNCH(k_pd_flush, A, elem, gslot, A, sa_final);
}
void pd_startflags(const u32* sa_final, const u32* isstart, const u32* wstart, u32 W, u32* flag, u32* val, cudaStream_t st) {
  PD_LAUNCH(k_pd_startflags, W, sa_final, isstart, wstart, W, flag, val);
}
void pd_emit(const u32* val, const u32* dst, u32 W, u32* out, cudaStream_t st) { PD_LAUNCH(k_pd_emit, W, val, dst, W, out); }
void pd_key2_plain(const u32* elem, const u32* rank, u32 A, u32 W, u32 off, u32* key2, cudaStream_t st) {
  PD_LAUNCH(k_pd_key2_plain, A, elem, rank, A, W, off, key2);
}
void pd_seg_small(const u32* headslot, u32 ngroups, u32 A, u32* elem, u32* key2, u32* lists, u32 list_cap, u32* counters,
                  cudaStream_t st) {
  PD_LAUNCH(k_pd_seg_small, ngroups, headslot, ngroups, A, elem, key2, lists, list_cap, counters);
}
template <int NW, int R>
static void pd_seg_class(const u32* list, u32 cnt, const u32* headslot, u32 ngroups, u32 A, u32* elem, u32* key2,
                         cudaStream_t st) {
  if (!cnt) return;
  u32 grid = NW == 1 ? (cnt + 7) / 8 : cnt;
  if (grid > 524288u) grid = 524288u;
  k_pd_seg_team<NW, R><<<grid, NW == 1 ? 256 : NW * 32, 0, st>>>(list, cnt, headslot, ngroups, A, elem, key2);
  CK_LAUNCH(); g_launches++;
}
void pd_seg_teams(const u32* lists, u32 list_cap, const u32* cn, const u32* headslot, u32 ngroups, u32 A, u32* elem, u32* key2,
                  cudaStream_t st) {
  pd_seg_class<1, 1>(lists + 0ull * list_cap, cn[0], headslot, ngroups, A, elem, key2, st);
  pd_seg_class<1, 2>(lists + 1ull * list_cap, cn[1], headslot, ngroups, A, elem, key2, st);
  pd_seg_class<1, 4>(lists + 2ull * list_cap, cn[2], headslot, ngroups, A, elem, key2, st);
  pd_seg_class<1, 8>(lists + 3ull * list_cap, cn[3], headslot, ngroups, A, elem, key2, st);
  pd_seg_class<8, 2>(lists + 4ull * list_cap, cn[4], headslot, ngroups, A, elem, key2, st);
  pd_seg_class<16, 2>(lists + 5ull * list_cap, cn[5], headslot, ngroups, A, elem, key2, st);
  pd_seg_class<16, 4>(lists + 6ull * list_cap, cn[6], headslot, ngroups, A, elem, key2, st);
}
void pd_heads_seg(const u32* key2, u32 A, u32* head, u32* isnew, cudaStream_t st) { PD_LAUNCH(k_pd_heads_seg, A, key2, A, head, isnew); }
void pd_big_sizes(const u32* biglist, u32 nb, const u32* headslot, u32 ngroups, u32 A, u32* bsize, cudaStream_t st) {
  PD_LAUNCH(k_pd_big_sizes, nb, biglist, nb, headslot, ngroups, A, bsize);
}
void pd_big_fill(const u32* biglist, const u32* boff, u32 nb, u32 B, const u32* headslot, const u32* elem, const u32* key2,
                 int rb, u64* comp, u32* val, u32* bslot, cudaStream_t st) {
  PD_LAUNCH(k_pd_big_fill, B, biglist, boff, nb, B, headslot, elem, key2, rb, comp, val, bslot);
}
void pd_big_writeback(const u64* comp, const u32* val, const u32* bslot, u32 B, int rb, u32* elem, u32* key2, cudaStream_t st) {
  PD_LAUNCH(k_pd_big_writeback, B, comp, val, bslot, B, rb, elem, key2);
}

}  // namespace kiss
