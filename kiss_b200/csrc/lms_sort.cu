// lms_sort.cu — K3: the k-ordered LMS sort beyond the first radix key.
//
// Replaces lms_suffix_direct_sort_dna (kiss1_core.hpp:23-145).  The reference's order is:
// the first D(k) = 125 * (k/125 + 1) bases, then the text position — the comparator lambda
// (kiss1_core.hpp:94-135) compares whole 125-base blocks while sorted_len <= k.  For suffixes with fewer than
// D(k) bases left the lambda's depth shrinks; those few ("short" suffixes, at most D/2 of them) are ordered
// with a literal restatement of the lambda and inserted into the sorted long ones.
//
// After the stable LSD radix sort on the first 16 bases (radix.cu) equal keys are adjacent and in ascending
// text position.  Groups of equal keys are finished by size:
//   tiny   (<= 8)     k_groups_tiny: one thread per group, insertion sort on (next 32 bases, deeper words, position)
//   medium (<= 2048)  k_groups_medium: one team (a warp up to 256 elements, a block above) runs PIVOT ROUNDS — every
//                     element is compared with its segment's pivot (block teams: the base-wise majority of 3-5 copies)
//                     and gets one key [segment | side | first differing base | next 14 bases | slot]; the first round of
//                     a block team is ordered by a counting sort on (side, q, base) + in-bucket ranking, later rounds by
//                     ranking inside the short segments or the bitonic network of team.cuh; segments of <= 4 are finished
//                     by one thread each
//   big               the same pivot keys in global memory (k_big_fill), ordered with the LSD radix sort; the resulting
//                     classes go through the same three cases at the next level
// All paths are stable with respect to text position, which is the reference's final tie-break (:131-133).
// The unbounded order runs under a work cap (SortParams::abort) and falls back to rank doubling (pd.cu, api.cu).
#include <cstdlib>
#include "common.cuh"
#include "kernels.h"
#include "team.cuh"

namespace kiss {

// true iff suffix p sorts before suffix q; both are known equal on their first h bases.
// Bounded mode: both suffixes have at least D bases.  Unbounded (D = KISS_UNBOUNDED): the shorter suffix
// first when one is a prefix of the other (kiss1_core.hpp:134).
// the work cap of the unbounded pivot rounds: `it` = words compared so far by this comparison
__device__ __forceinline__ bool over_cap(const SortParams& sp, u32 it) {
  if (!sp.abort) return false;
  if (it >= sp.cap_words) { *sp.abort = 1u; return true; }
  return (it & 31u) == 31u && *reinterpret_cast<volatile u32*>(sp.abort) != 0u;
}

__device__ __forceinline__ bool suffix_less(const SortParams& sp, u32 p, u32 q, u64 h) {
  u32 it = 0;
  for (u64 off = h; off < sp.D; off += 32) {
    if (over_cap(sp, it++)) return p < q;  // (the result is discarded)
    if ((u64)p + off >= sp.n || (u64)q + off >= sp.n) return p > q;
    u64 a = word_at(sp.T, (u64)p + off), b = word_at(sp.T, (u64)q + off);
    u64 rem = sp.D - off;
    if (rem < 32) {
      u64 mk = ~0ull << (64 - 2 * rem);
      a &= mk;
      b &= mk;
    }
    if (a != b) return a < b;
  }
  return p < q;
}

// ------------------------------------------------------------------------------------------- tiny groups
// One block looks at 2048 consecutive sorted keys (+ a halo of TINY_MAX).  Three dense phases:
//  1. every thread compares its 8 consecutive keys (two 16-byte loads, neighbours by shuffle) and queues the
//     tile-local index of every tied element and of every group head in shared memory;
//  2. the tied elements (14 % on uniform DNA) fetch position and next 32 bases — the queue makes this phase run
//     with full warps instead of one active lane in seven;
//  3. one thread per group head orders its group of <= TINY_MAX by (next word, deeper words, position).
#define TG_THREADS 256
#define TG_ITEMS 8
#define TG_TILE (TG_THREADS * TG_ITEMS)
template <typename K>
__global__ void __launch_bounds__(TG_THREADS) k_groups_tiny(const K* __restrict__ keys, u32* __restrict__ pos, u32 N,
                                                           SortParams sp, u64 h, GroupLists gl) {
  __shared__ u32 spos[TG_TILE + TINY_MAX];
  __shared__ u64 sw[TG_TILE + TINY_MAX];
  __shared__ unsigned short stied[TG_TILE + TINY_MAX];  // tile-local indices of the tied elements
  __shared__ unsigned short sheads[TG_TILE / 2];        // ... of the group heads (a group has >= 2 elements)
  __shared__ u32 s_tied, s_nheads, s_ntied;
  const u32 tid = threadIdx.x, lane = tid & 31u;
  if (tid == 0) { s_tied = 0; s_nheads = 0; s_ntied = 0; }
  __syncthreads();
  const u32 base = blockIdx.x * TG_TILE;  // (N < 2^32)
  const bool deep = h < sp.D;
  const u64 rem = sp.D - h;
  {
    // ---- phase 1: k[i + 1] = key of element t0 + i, k[0] / k[9] the neighbours
    const u32 t0 = tid * TG_ITEMS, e0 = base + t0;
    K k[TG_ITEMS + 2];
    if (e0 + TG_ITEMS <= N && ((uintptr_t)keys & 15) == 0) {
      constexpr int PER = 16 / sizeof(K);
      const uint4* src = reinterpret_cast<const uint4*>(keys + e0);
#pragma unroll
      for (int q = 0; q < TG_ITEMS / PER; q++) {
        uint4 v = __ldg(src + q);
        if (sizeof(K) == 4) {
          k[1 + 4 * q] = (K)v.x; k[2 + 4 * q] = (K)v.y; k[3 + 4 * q] = (K)v.z; k[4 + 4 * q] = (K)v.w;
        } else {
          k[1 + 2 * q] = (K)(((u64)v.y << 32) | v.x); k[2 + 2 * q] = (K)(((u64)v.w << 32) | v.z);
        }
      }
    } else {
#pragma unroll
      for (int i = 0; i < TG_ITEMS; i++) k[1 + i] = e0 + i < N ? keys[e0 + i] : (K)0;
    }
    K up = __shfl_up_sync(0xffffffffu, k[TG_ITEMS], 1), dn = __shfl_down_sync(0xffffffffu, k[1], 1);
    k[0] = lane > 0 ? up : (e0 > 0 && e0 <= N ? keys[e0 - 1] : (K)0);
    k[TG_ITEMS + 1] = lane < 31 ? dn : (e0 + TG_ITEMS < N ? keys[e0 + TG_ITEMS] : (K)0);
    u32 tmask = 0, hmask = 0;
#pragma unroll
    for (int i = 0; i < TG_ITEMS; i++) {
      const u32 e = e0 + i;
      const bool prev = e > 0 && k[i] == k[i + 1], next = e + 1 < N && k[i + 2] == k[i + 1];
      if (e < N && (prev || next)) tmask |= 1u << i;
      if (e < N && next && !prev) hmask |= 1u << i;
    }
    // a thread whose 8 keys and both neighbours are all equal sits inside a run of >= 10 > TINY_MAX: that group is
    // not handled here (phase 3 hands it on by its head), so its elements need no position / text gathers
    // (repeat-rich text: most tied elements are in such runs)
    if (e0 > 0 && e0 + TG_ITEMS < N && k[0] == k[1] && k[TG_ITEMS + 1] == k[TG_ITEMS]) {
      bool alleq = true;
#pragma unroll
      for (int i = 1; i < TG_ITEMS; i++) alleq &= k[i] == k[i + 1];
      if (alleq) tmask = 0;
    }
    // the halo: elements past the tile that may belong to a group starting inside it
    bool halo_tied = false;
    if (tid < TINY_MAX) {
      const u32 e = base + TG_TILE + tid;
      if (e < N) {
        K kk = keys[e];
        halo_tied = keys[e - 1] == kk || (e + 1 < N && keys[e + 1] == kk);
      }
    }
    // queue slots: one atomic per warp and queue
    u32 cnt = (u32)__popc(tmask) + (halo_tied ? 1u : 0u) + ((u32)__popc(hmask) << 16);
    u32 incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      u32 x = __shfl_up_sync(0xffffffffu, incl, o);
      if ((int)lane >= o) incl += x;
    }
    u32 wbase = 0;
    if (lane == 31) {
      u32 a = atomicAdd(&s_ntied, incl & 0xFFFFu), b = atomicAdd(&s_nheads, incl >> 16);
      wbase = a | (b << 16);
    }
    wbase = __shfl_sync(0xffffffffu, wbase, 31);
    u32 ex = incl - cnt;
    u32 ot = (wbase & 0xFFFFu) + (ex & 0xFFFFu), oh = (wbase >> 16) + (ex >> 16);
#pragma unroll
    for (int i = 0; i < TG_ITEMS; i++) {
      if ((tmask >> i) & 1u) stied[ot++] = (unsigned short)(t0 + i);
      if ((hmask >> i) & 1u) sheads[oh++] = (unsigned short)(t0 + i);
    }
    if (halo_tied) stied[ot] = (unsigned short)(TG_TILE + tid);
  }
  __syncthreads();
  // ---- phase 2: dense gathers
  if (deep) {
    // (four elements per thread and step: the four position loads, then the eight text loads, are in flight together —
    // one element per step left two dependent DRAM latencies exposed between the barriers)
    const u32 ntied = s_ntied;
    for (u32 j0 = 0; j0 < ntied; j0 += 4 * TG_THREADS) {
      u32 t[4], pp[4];
#pragma unroll
      for (int q = 0; q < 4; q++) {
        const u32 j = j0 + q * TG_THREADS + tid;
        t[q] = j < ntied ? stied[j] : 0xFFFFFFFFu;
        pp[q] = j < ntied ? pos[base + t[q]] : 0u;
      }
      u64 wv[4];
#pragma unroll
      for (int q = 0; q < 4; q++) wv[q] = t[q] != 0xFFFFFFFFu ? word_at(sp.T, (u64)pp[q] + h) : 0ull;
#pragma unroll
      for (int q = 0; q < 4; q++) {
        if (t[q] != 0xFFFFFFFFu) {
          if (rem < 32) wv[q] &= ~0ull << (64 - 2 * rem);
          spos[t[q]] = pp[q];
          sw[t[q]] = wv[q];
        }
      }
    }
  }
  __syncthreads();
  // ---- phase 3: one thread per group
  const u32 nheads = s_nheads;
  for (u32 hi = tid; hi < nheads; hi += TG_THREADS) {
    const u32 t = sheads[hi];
    const u32 e = base + t;
    const K kk = keys[e];
    u32 g = 2;
    while (g <= TINY_MAX && e + g < N && keys[e + g] == kk) g++;
    if (g > TINY_MAX) {
      u32 idx = atomicAdd(&gl.counters[0], 1u);
      if (idx < gl.gcap) gl.glist[idx] = e; else gl.counters[3] = 1;
      continue;
    }
    atomicAdd(&s_tied, g);
    if (!deep) continue;  // fully compared: the stable order (ascending position) is final
    // near the end of the text (unbounded mode) or of the compared depth: plain comparator on positions
    bool slow = false;
    for (u32 a = 0; a < g; a++) slow |= ((u64)spos[t + a] + h + 32 > sp.n);
    bool changed = false;
    for (u32 a = 1; a < g; a++) {
      u32 xp = spos[t + a];
      u64 xw = sw[t + a];
      u32 b = a;
      while (b > 0) {
        u32 yp = spos[t + b - 1];
        u64 yw = sw[t + b - 1];
        bool less = slow ? suffix_less(sp, xp, yp, h) : (xw != yw ? xw < yw : suffix_less(sp, xp, yp, h + 32));
        if (!less) break;
        spos[t + b] = yp;
        sw[t + b] = yw;
        b--;
      }
      if (b != a) {
        spos[t + b] = xp;
        sw[t + b] = xw;
        changed = true;
      }
    }
    if (changed)
      for (u32 a = 0; a < g; a++) pos[e + a] = spos[t + a];
  }
  __syncthreads();
  if (tid == 0 && s_tied) atomicAdd(gl.tied, (u64)s_tied);
}

// depth (bases) the elements of a group of equal composite keys are known to share; `kind`: 0 = plain keys (the
// level's depth h), 1 / 2 = pivot keys of the big path in the bounded / unbounded layout (see k_big_fill)
__device__ __forceinline__ u32 group_depth(u64 key, int kind, u32 h) {
  if (kind == 1) {
    const u64 side = (key >> 40) & 3ull, qk = (key >> 28) & 0xFFFull;
    return (u32)((side == 2 ? 4095ull - qk : qk) + 14ull);
  }
  if (kind == 2) {
    const u64 side = (key >> 35) & 3ull, qk = (key >> 3) & 0xFFFFFFFFull;
    const u64 d = (side == 2 ? 0xFFFFFFFFull - qk : qk) + 1ull;
    return (u32)(d < 0xFFFFFFFFull ? d : 0xFFFFFFFFull);
  }
  return h;
}

template <typename K>
__global__ void __launch_bounds__(128) k_groups_classify(const K* __restrict__ keys, u32 N, u32 ng, GroupLists gl,
                                                        int kind, u32 h) {
  const u32 g = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31u;
  const bool valid = g < ng;
  u64 start = 0;
  u32 size = 0;
  K k = 0;
  if (valid) {
    start = gl.glist[g];
    k = keys[start];
    // keys[start + TINY_MAX] == k is known; gallop, then bisect for the first index whose key differs
    u64 lo = start + TINY_MAX, hi, step = TINY_MAX;
    for (;;) {
      u64 nx = lo + step;
      if (nx >= N) { hi = N; break; }
      if (keys[nx] != k) { hi = nx; break; }
      lo = nx;
      step <<= 1;
    }
    while (hi - lo > 1) {
      u64 mid = lo + (hi - lo) / 2;
      if (keys[mid] == k) lo = mid; else hi = mid;
    }
    size = (u32)(hi - start);
  }
  // statistics and list slots: one atomic per warp and class (millions of groups on repeat-rich text)
  const u32 wsum = __reduce_add_sync(0xffffffffu, size);
  if (lane == 0 && wsum) atomicAdd(gl.tied, (u64)wsum);
  const u32 cls = !valid ? 7u : size <= MED_MAX ? med_class_of(size) : 6u;
#pragma unroll
  for (u32 c = 0; c < MED_CLASSES; c++) {
    const u32 mask = __ballot_sync(0xffffffffu, cls == c);
    if (cls == c) {
      const u32 leader = __ffs(mask) - 1u;
      const u32 esum = __reduce_add_sync(mask, size);
      u32 base = 0;
      if (lane == leader) {
        base = atomicAdd(&gl.counters[4 + c], (u32)__popc(mask));
        atomicAdd(&gl.counters[10 + c], esum);
      }
      base = __shfl_sync(mask, base, leader);
      const u64 j = med_class_off(gl.medN, c) + base + __popc(mask & ((1u << lane) - 1u));
      gl.med[2 * j] = (u32)start;
      gl.med[2 * j + 1] = size;
    }
  }
  if (cls == 6u) {
    u32 j = atomicAdd(&gl.counters[2], 1u);
    gl.big[3 * j] = (u32)start;
    gl.big[3 * j + 1] = size;
    gl.big[3 * j + 2] = group_depth((u64)k, kind, h);
  }
}

void groups_tiny_u32(const u32* keys, u32* pos, u32 N, SortParams sp, u64 h, GroupLists gl, cudaStream_t st) {
  if (!N) return;
  ProfScope ps(PK_GROUPS_TINY, st);
  if (ps.p) ps.p->bytes[PK_GROUPS_TINY] += 8ull * N;  // key + position
  k_groups_tiny<u32><<<(N + TG_TILE - 1) / TG_TILE, TG_THREADS, 0, st>>>(keys, pos, N, sp, h, gl);
  CK_LAUNCH(); g_launches++;
}
void groups_tiny_u64(const u64* keys, u32* pos, u32 N, SortParams sp, u64 h, GroupLists gl, cudaStream_t st) {
  if (!N) return;
  ProfScope ps(PK_GROUPS_TINY, st);
  if (ps.p) ps.p->bytes[PK_GROUPS_TINY] += 12ull * N;
  k_groups_tiny<u64><<<(N + TG_TILE - 1) / TG_TILE, TG_THREADS, 0, st>>>(keys, pos, N, sp, h, gl);
  CK_LAUNCH(); g_launches++;
}
void groups_classify_u32(const u32* keys, u32 N, u32 ng, GroupLists gl, u32 h, cudaStream_t st) {
  if (!ng) return;
  ProfScope ps(PK_GROUPS_OTHER, st);
  k_groups_classify<u32><<<(ng + 127) / 128, 128, 0, st>>>(keys, N, ng, gl, 0, h);
  CK_LAUNCH(); g_launches++;
}
void groups_classify_u64(const u64* keys, u32 N, u32 ng, GroupLists gl, int kind, u32 h, cudaStream_t st) {
  if (!ng) return;
  ProfScope ps(PK_GROUPS_OTHER, st);
  k_groups_classify<u64><<<(ng + 127) / 128, 128, 0, st>>>(keys, N, ng, gl, kind, h);
  CK_LAUNCH(); g_launches++;
}

// ------------------------------------------------------------------------------------------- medium groups
// One team (a warp up to 256 elements, a block above) refines one group by PIVOT ROUNDS.  In a round every element
// that still shares its segment with another is compared with the segment's pivot from the segment's depth on: it
// finds the first base q where it differs (word-wise, independent loads), which side of the pivot it falls on, and
// its own next bases from q.  Its key
//     hi = [segment:11 | side:2 | q (descending on the greater side):12 | 14 bases:28 | slot:11]   lo = the next 32 bases
// then orders the whole segment: before the pivot by ascending q, after it by descending q, equal (q, bases) classes
// become new segments at depth q + 46, and everything equal to the pivot up to D is final (the slot field keeps equal
// keys in their current order, which is ascending text position).  Long common prefixes of repeat copies are skipped
// in one step, and a group without common prefix is split by 46 bases per round (the first version carried 14 bases
// in a single word: copies that diverge by 5-15 % needed ~7 rounds of the sorting network per group, 7700 thread
// instructions per element in the largest class; profiles/r2_a_bench_*.json).
// The pivot of a block team is VIRTUAL: the base-wise majority of three suffixes of the segment.  Any string works as a
// pivot; a real element carries its own mutations — every other copy differs from it THERE, so the main lineage of a
// group of near-identical copies advanced only to the pivot's next mutation per round — while two of three copies agree
// almost everywhere, so nearly every element is split off by its OWN first mutation.  (Two of the three agreeing fixes
// both bits of a base; where all three differ the field is arbitrary, which is as good as any.  The three loads are
// the same for every thread of a segment.)  Warp teams keep the segment's first element: their groups are small and
// the extra loads cost more than the rounds they save.
// Unbounded depth (-k -1 / PREFIX_DOUBLING): q needs 32 bits, so the key is one word carrying one base, and a suffix
// that ends inside the comparison is smaller than one that goes on (kiss1_core.hpp:134).
// The keys live in registers — lane l of warp w holds the slots (w R + r) 32 + l, r < R — and are bitonic-sorted
// with warp shuffles (partner distance < 32), register swaps (< 32 R) and, in the block team only, shared memory.
#define MK_SIDE_SHIFT 51
#define MK_SEG_SHIFT 53
#define MK_EQUAL 1ull
#define MK_WIDE_BASES 46ull  // bases of its own an element of the bounded layout carries from q on (14 in hi, 32 in lo)
__device__ __forceinline__ u64 maj3(u64 a, u64 b, u64 c) { return (a & b) | ((a | b) & c); }
// bit-wise majority of five (a bit-sliced adder: sum = s1 + 2 (c0 + c1) >= 3)
__device__ __forceinline__ u64 maj5(u64 a, u64 b, u64 c, u64 d, u64 e) {
  const u64 s0 = a ^ b ^ c, c0 = maj3(a, b, c), s1 = s0 ^ d ^ e, c1 = maj3(s0, d, e);
  return (c0 & c1) | ((c0 ^ c1) & s1);
}

// bounded layout: key of suffix p against the pivot, whose 32 bases from depth q on are pivot_word(q) (q runs over
// d, d + 32, ...); `d` = depth all are known equal to.  Returns hi (without segment and slot); lo = the 32 bases after
// the 14 in hi, cut at D.
template <typename PW>
__device__ __forceinline__ u64 pivot_key_pw(const SortParams& sp, u32 p, u64 d, u64& lo, PW pivot_word) {
  lo = 0;
  u64 q = d;
  while (q < sp.D) {
    u64 wa = word_at(sp.T, (u64)p + q), wb = pivot_word(q);
    const u64 rem = sp.D - q;
    if (rem < 32) { const u64 mk = ~0ull << (64 - 2 * rem); wa &= mk; wb &= mk; }
    const u64 x = wa ^ wb;
    if (x) {
      q += (u64)(__clzll((long long)x) >> 1);
      const u64 side = wa > wb ? 2ull : 0ull;
      const u64 w0 = word_at(sp.T, (u64)p + q), w1 = word_at(sp.T, (u64)p + q + 32);
      u64 b14 = w0 >> 36;
      const u64 rem2 = sp.D - q;  // >= 1
      if (rem2 < 14) b14 &= (~0ull << (28 - 2 * rem2)) & 0xFFFFFFFull;
      if (rem2 > 14) {
        lo = (w0 << 28) | (w1 >> 36);
        if (rem2 < MK_WIDE_BASES) lo &= ~0ull << (64 - 2 * (rem2 - 14));
      }
      const u64 qk = side == 2 ? 4095ull - q : q;
      return (side << MK_SIDE_SHIFT) | (qk << 39) | (b14 << MK_SLOT_BITS);
    }
    q += 32;
  }
  return MK_EQUAL << MK_SIDE_SHIFT;
}
// ... against p1 alone, or against the majority of p1, p2, p3
__device__ __forceinline__ u64 pivot_key_b(const SortParams& sp, u32 p, u32 p1, u32 p2, u32 p3, bool maj, u64 d, u64& lo) {
  return pivot_key_pw(sp, p, d, lo, [&](u64 q) {
    u64 wb = word_at(sp.T, (u64)p1 + q);
    if (maj) wb = maj3(wb, word_at(sp.T, (u64)p2 + q), word_at(sp.T, (u64)p3 + q));
    return wb;
  });
}
// unbounded layout: key of suffix p against the real pivot ph
__device__ __forceinline__ u64 pivot_key_u(const SortParams& sp, u32 p, u32 ph, u64 d) {
  u64 q = d, side = MK_EQUAL;
  const u64 lp = sp.n - p, lh = sp.n - ph, lim = lp < lh ? lp : lh;  // bases both suffixes have
  const u64 L = lim < sp.D ? lim : sp.D;
  u32 it = 0;
  while (q < L) {
    if (over_cap(sp, it++)) return MK_EQUAL << MK_SIDE_SHIFT;  // (the result is discarded)
    u64 wa = word_at(sp.T, (u64)p + q), wb = word_at(sp.T, (u64)ph + q);
    const u64 rem = L - q;
    if (rem < 32) { const u64 mk = ~0ull << (64 - 2 * rem); wa &= mk; wb &= mk; }
    const u64 x = wa ^ wb;
    if (x) {
      q += (u64)(__clzll((long long)x) >> 1);
      side = wa > wb ? 2ull : 0ull;
      break;
    }
    q += 32;
  }
  u64 b = 0;  // 0 = the suffix ended at q
  if (side == MK_EQUAL) {
    if (L == sp.D) return MK_EQUAL << MK_SIDE_SHIFT;  // equal on the whole compared depth: position decides
    q = lim;  // one is a prefix of the other: the shorter first
    side = lp < lh ? 0ull : 2ull;
  }
  if ((u64)p + q < sp.n) b = 1ull + (word_at(sp.T, (u64)p + q) >> 62);
  const u64 qk = side == 2 ? 0xFFFFFFFFull - q : q;
  return (side << MK_SIDE_SHIFT) | (qk << 19) | (b << 16);
}
// depth at which the elements of a class (equal keys above the slot field) are known equal; `own` = bases of its own
// the key carries from q on (14 for the one-word keys of the big path, MK_WIDE_BASES in the teams)
template <bool UNB>
__device__ __forceinline__ u64 class_depth(u64 key, u64 own) {
  const u64 side = (key >> MK_SIDE_SHIFT) & 3ull;
  if (!UNB) {
    const u64 qk = (key >> 39) & 0xFFFull;
    return (side == 2 ? 4095ull - qk : qk) + own;
  }
  const u64 qk = (key >> 19) & 0xFFFFFFFFull;
  return (side == 2 ? 0xFFFFFFFFull - qk : qk) + 1ull;
}

#define MED_THREADS(NW) ((NW) == 1 ? 256 : (NW) * 32)
#define MED_NBK 4096u      // bucket counters of the first round of a block team
#define MED_VCACHE 64u     // words of the virtual pivot kept in shared memory in the first round
template <int NW, int R, bool UNB, bool WIDE>
__global__ void __launch_bounds__(MED_THREADS(NW), (R >= 8 ? 2 : (NW > 8 ? 2 : 4))) k_groups_medium(u32* __restrict__ pos, const u32* __restrict__ med,
                                                                        u32 nmed, SortParams sp, u64 h, u32 rank_scale, u32 vpivot,
                                                                        u32 bucket_cost, u32 fin_max, u64* __restrict__ dbg) {
  static_assert(!(UNB && WIDE), "the unbounded layout has one-word keys");
  constexpr u32 P = NW * R * 32;
  constexpr u32 TEAMS = NW == 1 ? 8 : 1;
  static_assert(P <= 2048, "slot field is 11 bits");
  // dynamic shared memory: key words (hi, and lo when WIDE), positions, segment heads, depths; block teams: bucket
  // counters and the cached pivot words of the first round
  extern __shared__ __align__(16) unsigned char med_smem[];
  u64* s_keyh = reinterpret_cast<u64*>(med_smem);
  u64* s_keyl = s_keyh + TEAMS * P;  // (WIDE only)
  u32* s_pos = reinterpret_cast<u32*>(s_keyh + (WIDE ? 2 : 1) * TEAMS * P);
  u32* s_seg = s_pos + TEAMS * P;
  u32* s_dep = s_seg + TEAMS * P;  // depth of the segment, kept at its head slot
  u64* s_vp = reinterpret_cast<u64*>(s_dep + TEAMS * P);  // (NW > 1) [MED_VCACHE]
  u32* s_cnt = reinterpret_cast<u32*>(s_vp + MED_VCACHE);  // (NW > 1) [MED_NBK]
  __shared__ u32 s_ws[33];
  __shared__ u32 s_scan[NW];
  __shared__ u32 s_qn[TEAMS];  // small segments queued for their finishing threads
  const u32 lane = threadIdx.x & 31u, wb = threadIdx.x >> 5;
  const u32 team = NW == 1 ? wb : 0u, w = NW == 1 ? 0u : wb;
  u64* skh = s_keyh + team * P;
  u64* skl = s_keyl + team * P;
  u32* spos = s_pos + team * P;
  u32* sseg = s_seg + team * P;
  u32* sdep = s_dep + team * P;
  const bool maj = !UNB && vpivot != 0;
  for (u32 gi = blockIdx.x * TEAMS + team; gi < nmed; gi += gridDim.x * TEAMS) {
    const u32 start = med[2 * gi], size = med[2 * gi + 1];
#pragma unroll
    for (int r = 0; r < R; r++) {
      const u32 i = ((w * R + r) << 5) | lane;
      spos[i] = i < size ? pos[start + i] : KISS_EMPTY;
      sseg[i] = i < size ? 0u : 0xFFFFFFFFu;
    }
    if (threadIdx.x % (NW * 32) == 0) { sdep[0] = (u32)h; sdep[1] = size; s_qn[team] = 0; }
    team_sync<NW>();
    // Two ways to order the segments of a round.  The bitonic network costs the same whatever the segments look
    // like (LOGP (LOGP + 1) / 2 steps of ~12 instructions per key); counting, for every element, the smaller keys
    // of its own segment costs ~4 instructions per segment member.  After the first round the segments are short,
    // so the later rounds rank.
    constexpr u32 LOGP = P == 32 ? 5 : P == 64 ? 6 : P == 128 ? 7 : P == 256 ? 8 : P == 512 ? 9 : P == 1024 ? 10 : 11;
    const u32 RANK_MAX = rank_scale * LOGP * (LOGP + 1) / 2;  // longest segment for which ranking is the cheaper way
    u32 maxlen = size;                               // longest segment with >= 2 elements (team-uniform)
    u32 dbg_round = 0;
    bool first = true;
    // first round of a block team: the virtual pivot's words over [h, D) are computed once per group
    u32 nvp = 0;
    if constexpr (NW > 1 && !UNB) {
      if (maj) {
        const u64 span = sp.D > h ? sp.D - h : 0;
        nvp = (u32)min((u64)MED_VCACHE, (span + 31) / 32);
        if (threadIdx.x < nvp) {
          const u64 q = h + 32ull * threadIdx.x;
          // (computed once per group, so five samples cost nothing: the majority is then off the consensus — which sends
          // the whole group into ONE bucket at that base — where three of five copies share a mutation, not two of three)
          s_vp[threadIdx.x] = maj5(word_at(sp.T, (u64)spos[0] + q), word_at(sp.T, (u64)spos[size >> 2] + q),
                                   word_at(sp.T, (u64)spos[size >> 1] + q), word_at(sp.T, (u64)spos[size - 1 - (size >> 2)] + q),
                                   word_at(sp.T, (u64)spos[size - 1] + q));
        }
        __syncthreads();
      }
    }
    for (;;) {
      u64 xh[R], xl[R];
      bool anyact = false, moved = false;
      u32 actm = 0;
#pragma unroll
      for (int r = 0; r < R; r++) {
        const u32 i = ((w * R + r) << 5) | lane;
        u64 key = ~0ull, klo = ~0ull;
        if (i < size) {
          const u32 sg = sseg[i];
          const bool act = sg != i || (i + 1 < size && sseg[i + 1] == sg);
          key = 0;
          klo = 0;
          if (act) {
            anyact = true;
            actm |= 1u << r;
            key = MK_EQUAL << MK_SIDE_SHIFT;
            if constexpr (UNB) {
              if (sg != i) key = pivot_key_u(sp, spos[i], spos[sg], sdep[sg]);
            } else if (maj) {
              const u32 end = sdep[sg + 1];  // (a segment that is still refined keeps its end in the slot after its head)
              const u32 p1 = spos[sg], p2 = spos[sg + ((end - sg) >> 1)], p3 = spos[end - 1];
              if (NW > 1 && first) {
                key = pivot_key_pw(sp, spos[i], h, klo, [&](u64 q) {
                  const u64 k = (q - h) >> 5;
                  if (k < nvp) return s_vp[k];
                  return maj5(word_at(sp.T, (u64)p1 + q), word_at(sp.T, (u64)spos[size >> 2] + q), word_at(sp.T, (u64)p2 + q),
                              word_at(sp.T, (u64)spos[size - 1 - (size >> 2)] + q), word_at(sp.T, (u64)p3 + q));
                });
              } else {
                key = pivot_key_b(sp, spos[i], p1, p2, p3, true, sdep[sg], klo);
              }
            } else if (sg != i) {
              key = pivot_key_b(sp, spos[i], spos[sg], 0u, 0u, false, sdep[sg], klo);
            }
            moved |= ((key >> MK_SIDE_SHIFT) & 3ull) != MK_EQUAL;
          }
          key |= ((u64)sg << MK_SEG_SHIFT) | i;
        }
        xh[r] = key;
        xl[r] = klo;
      }
      if (!team_or<NW>(anyact)) break;
      if (dbg) {  // KISS_B200_DEBUG_GROUPS: rounds, active elements and longest segment per round
        const u32 na = __reduce_add_sync(0xffffffffu, (u32)__popc(actm));
        if (lane == 0) atomicAdd(&dbg[16 + min(dbg_round, 15u)], (u64)na);
        if (threadIdx.x % (NW * 32) == 0) {
          atomicAdd(&dbg[0], 1ull);
          atomicAdd(&dbg[maxlen > RANK_MAX ? 1 : 2], 1ull);
          atomicAdd(&dbg[32 + min(dbg_round, 15u)], (u64)maxlen);
          atomicAdd(&dbg[48 + min(dbg_round, 15u)], 1ull);
        }
        dbg_round++;
      }
      if (team_or<NW>(moved)) {
        bool sorted = false;
        if constexpr (NW > 1 && !UNB) {
          if (first && bucket_cost && maxlen > RANK_MAX) {
            // First round of a block team (one segment = the whole group): the (side, q, first base) part of the key
            // alone spreads the elements over up to MED_NBK buckets — a counting sort with shared-memory atomics,
            // unordered inside a bucket — and every element then counts the smaller keys of its own bucket.  ~12
            // instructions per bucket member against ~1750 per element for the network (66 steps at 2048 slots on
            // two-word keys: 58 % of this kernel's instructions, profiles/r2_medium2048_wide.txt).  A group that
            // crowds one bucket (exact copies, a shared mutation) takes the network instead.
            const u32 span = (u32)(sp.D - h);  // bucket index: monotone in (side, q, base), quantised to fit
            u32 qs = 0;
            while (((4u * (2u * span + 1u)) >> qs) > MED_NBK) qs++;
            auto bucket_of = [&](u64 kh) {
              const u32 side = (u32)(kh >> MK_SIDE_SHIFT) & 3u, qk = (u32)(kh >> 39) & 0xFFFu, c = (u32)(kh >> 37) & 3u;
              const u32 q = side == 2 ? 4095u - qk : qk;
              const u32 i0 = side == 0 ? q - (u32)h : side == 1 ? span : span + 1u + ((u32)sp.D - 1u - q);
              return (4u * i0 + (side == 1 ? 0u : c)) >> qs;
            };
            for (u32 t = threadIdx.x; t < MED_NBK; t += NW * 32) s_cnt[t] = 0;
            __syncthreads();
#pragma unroll
            for (int r = 0; r < R; r++)
              if ((((w * R + r) << 5) | lane) < size) atomicAdd(&s_cnt[bucket_of(xh[r])], 1u);
            __syncthreads();
            {
              constexpr u32 PER = MED_NBK / (NW * 32);
              u32 c[PER], sum = 0, mx = 0, sq = 0;
#pragma unroll
              for (u32 e = 0; e < PER; e++) {
                c[e] = s_cnt[threadIdx.x * PER + e];
                mx = max(mx, c[e]);
                sum += c[e];
                sq += c[e] * c[e];
              }
              u32 tot;
              u32 ex = block_excl_scan(sum, s_ws, tot);
#pragma unroll
              for (u32 e = 0; e < PER; e++) {
                s_cnt[threadIdx.x * PER + e] = ex;
                ex += c[e];
              }
              // ranking costs one step per (element, member of its bucket) pair: sum of squares against size x the
              // network's ~bucket_cost steps per element
              sq = __reduce_add_sync(0xffffffffu, sq);
              if (lane == 0) s_scan[w] = sq;
              __syncthreads();
              u32 sqt = 0;
#pragma unroll
              for (u32 ww = 0; ww < NW; ww++) sqt += s_scan[ww];
              sorted = sqt <= bucket_cost * size;
              __syncthreads();  // (s_scan is reused below)
              if (dbg) {
                const u32 eqb = (4u * span) >> qs;  // the bucket of the elements equal to the pivot
                const u32 eqc = (eqb + 1 < MED_NBK ? s_cnt[eqb + 1] : size) - s_cnt[eqb];
                const u32 bmx = __reduce_max_sync(0xffffffffu, mx);
                if (lane == 0) atomicMax(&dbg[8], (u64)bmx);
                if (threadIdx.x == 0) {
                  atomicAdd(&dbg[sorted ? 3 : 4], 1ull);
                  if (!sorted && eqc * eqc > bucket_cost * size / 2) atomicAdd(&dbg[5], 1ull);
                  atomicAdd(&dbg[6], (u64)eqc);
                }
              }
            }
            if (sorted) {
#pragma unroll
              for (int r = 0; r < R; r++) {
                if ((((w * R + r) << 5) | lane) < size) {
                  const u32 t = atomicAdd(&s_cnt[bucket_of(xh[r])], 1u);
                  skh[t] = xh[r];
                  if constexpr (WIDE) skl[t] = xl[r];
                }
              }
              __syncthreads();  // s_cnt[b] is now the END of bucket b
              u32 pp[R], dst[R];
#pragma unroll
              for (int r = 0; r < R; r++) {
                const u32 t = ((w * R + r) << 5) | lane;
                pp[r] = KISS_EMPTY;
                dst[r] = t;
                xh[r] = ~0ull;
                xl[r] = ~0ull;
                if (t < size) {
                  xh[r] = skh[t];
                  if constexpr (WIDE) xl[r] = skl[t];
                  const u32 b = bucket_of(xh[r]);
                  const u32 lo_ = b ? s_cnt[b - 1] : 0u, hi_ = s_cnt[b];
                  u32 rank = 0;
                  for (u32 j = lo_; j < hi_; j++) {
                    u64 yl = 0;
                    if constexpr (WIDE) yl = skl[j];
                    rank += mk_less<WIDE>(skh[j], yl, xh[r], xl[r]) ? 1u : 0u;
                  }
                  dst[r] = lo_ + rank;
                  pp[r] = spos[(u32)xh[r] & 0x7FFu];
                }
              }
              __syncthreads();
#pragma unroll
              for (int r = 0; r < R; r++) {
                skh[dst[r]] = xh[r];
                if constexpr (WIDE) skl[dst[r]] = xl[r];
                spos[dst[r]] = pp[r];
              }
              __syncthreads();
#pragma unroll
              for (int r = 0; r < R; r++) {
                xh[r] = skh[((w * R + r) << 5) | lane];
                if constexpr (WIDE) xl[r] = skl[((w * R + r) << 5) | lane];
              }
              __syncthreads();
            }
          }
        }
        if (sorted) {
        } else if (maxlen > RANK_MAX) {
          team_sort<NW, R, WIDE>(xh, xl, w, lane, skh, skl);
          u32 pp[R];
#pragma unroll
          for (int r = 0; r < R; r++) pp[r] = xh[r] != ~0ull ? spos[(u32)xh[r] & 0x7FFu] : KISS_EMPTY;
          team_sync<NW>();
#pragma unroll
          for (int r = 0; r < R; r++) spos[((w * R + r) << 5) | lane] = pp[r];
        } else {
#pragma unroll
          for (int r = 0; r < R; r++) {
            skh[((w * R + r) << 5) | lane] = xh[r];
            if constexpr (WIDE) skl[((w * R + r) << 5) | lane] = xl[r];
          }
          team_sync<NW>();
          u32 pp[R], dst[R];
#pragma unroll
          for (int r = 0; r < R; r++) {
            const u32 i = ((w * R + r) << 5) | lane;
            pp[r] = spos[i];
            dst[r] = i;
            if ((actm >> r) & 1u) {
              const u32 sg = (u32)(xh[r] >> MK_SEG_SHIFT);  // head slot of the segment; its end is kept at sdep[head + 1]
              const u32 end = sdep[sg + 1];
              u32 rank = 0;
              for (u32 j = sg; j < end; j++) {
                u64 yl = 0;
                if constexpr (WIDE) yl = skl[j];
                rank += mk_less<WIDE>(skh[j], yl, xh[r], xl[r]) ? 1u : 0u;
              }
              dst[r] = sg + rank;
            }
          }
          team_sync<NW>();
#pragma unroll
          for (int r = 0; r < R; r++) {
            skh[dst[r]] = xh[r];
            if constexpr (WIDE) skl[dst[r]] = xl[r];
            spos[dst[r]] = pp[r];
          }
          team_sync<NW>();
#pragma unroll
          for (int r = 0; r < R; r++) {
            xh[r] = skh[((w * R + r) << 5) | lane];
            if constexpr (WIDE) xl[r] = skl[((w * R + r) << 5) | lane];
          }
          team_sync<NW>();
        }
      }
#pragma unroll
      for (int r = 0; r < R; r++) {
        skh[((w * R + r) << 5) | lane] = xh[r];
        if constexpr (WIDE) skl[((w * R + r) << 5) | lane] = xl[r];
      }
      team_sync<NW>();
      // new segment = running maximum of the head slots; everything equal to its pivot is final (own segment)
      u32 run = 0;  // head slot + 1 carried into the current row
      u32 sg[R];
      bool hdv[R];
#pragma unroll
      for (int r = 0; r < R; r++) {
        const u32 i = ((w * R + r) << 5) | lane;
        bool hd = false;
        if (i < size) {
          const u64 a = xh[r] >> MK_SLOT_BITS;
          hd = i == 0 || a != (skh[i - 1] >> MK_SLOT_BITS) || ((xh[r] >> MK_SIDE_SHIFT) & 3ull) == MK_EQUAL;
          if constexpr (WIDE) { if (!hd) hd = xl[r] != skl[i - 1]; }
        }
        hdv[r] = hd;
        u32 v = hd ? i + 1u : 0u;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          u32 u = __shfl_up_sync(0xffffffffu, v, o);
          if ((int)lane >= o) v = max(v, u);
        }
        v = max(v, run);
        sg[r] = v;
        run = __shfl_sync(0xffffffffu, v, 31);
      }
      u32 before = 0;
      if constexpr (NW > 1) {
        if (lane == 0) s_scan[w] = run;
        __syncthreads();
        for (u32 ww = 0; ww < w; ww++) before = max(before, s_scan[ww]);
      }
#pragma unroll
      for (int r = 0; r < R; r++) {
        const u32 i = ((w * R + r) << 5) | lane;
        if (i < size) {
          sseg[i] = max(sg[r], before) - 1u;
          if (hdv[r]) {
            const u64 dd = class_depth<UNB>(xh[r], WIDE ? MK_WIDE_BASES : 14ull);
            sdep[i] = (u32)(dd < 0xFFFFFFFFull ? dd : 0xFFFFFFFFull);
          }
        }
      }
      team_sync<NW>();
      // segment ends (kept in the slot after the head: a segment that is still refined has >= 2 elements) and the
      // longest segment
      // Segments of at most fin_max elements are FINISHED here by one thread each: an insertion sort with the word-wise
      // suffix comparison from the segment's depth on (stable: equal to D keeps text order).  After the first round
      // nearly all that is left are such crumbs (C3, largest class: 20 % of the elements, in segments of a handful),
      // and a team round costs the same barriers, scans and a full pass of the key loop whether 1400 or 14 elements
      // are still active (half of this kernel's instructions were key loops of later rounds with a few lanes alive,
      // profiles/r2_medium2048_narrow.txt).
      u32* queue = reinterpret_cast<u32*>(skh);  // (the key words are free until the next round)
      u32 ml = 0;
#pragma unroll
      for (int r = 0; r < R; r++) {
        const u32 i = ((w * R + r) << 5) | lane;
        if (i < size) {
          const u32 hd = sseg[i];
          if (hd != i && (i + 1 == size || sseg[i + 1] != hd)) {
            const u32 len = i + 1u - hd;
            if (len <= fin_max) {
              queue[atomicAdd(&s_qn[team], 1u)] = hd | (len << 16);
            } else {
              sdep[hd + 1] = i + 1u;
              ml = max(ml, len);
            }
          }
        }
      }
      team_sync<NW>();
      {
        const u32 nq = s_qn[team];
        for (u32 e = threadIdx.x % (NW * 32); e < nq; e += NW * 32) {
          const u32 hd = queue[e] & 0xFFFFu, len = queue[e] >> 16;
          const u64 dep = sdep[hd];
          for (u32 a = 1; a < len; a++) {
            const u32 x = spos[hd + a];
            u32 b = a;
            while (b > 0) {
              const u32 y = spos[hd + b - 1];
              if (!suffix_less(sp, x, y, dep)) break;
              spos[hd + b] = y;
              b--;
            }
            spos[hd + b] = x;
          }
          for (u32 a = 0; a < len; a++) sseg[hd + a] = hd + a;
        }
        team_sync<NW>();
        if (threadIdx.x % (NW * 32) == 0) s_qn[team] = 0;
      }
      ml = __reduce_max_sync(0xffffffffu, ml);
      if constexpr (NW > 1) {
        __syncthreads();  // (s_scan was read above)
        if (lane == 0) s_scan[w] = ml;
        __syncthreads();
#pragma unroll
        for (u32 ww = 0; ww < NW; ww++) ml = max(ml, s_scan[ww]);
      }
      maxlen = ml;
      first = false;
      team_sync<NW>();
    }
    team_sync<NW>();
#pragma unroll
    for (int r = 0; r < R; r++) {
      const u32 i = ((w * R + r) << 5) | lane;
      if (i < size) pos[start + i] = spos[i];
    }
    team_sync<NW>();
  }
}

u64* g_med_dbg = nullptr;  // KISS_B200_DEBUG_GROUPS: 6 classes x 64 counters (device), owned by api.cu

// KISS_B200_REAL_PIVOT=1: the first element of a segment as the pivot everywhere (the round-1 scheme, kept for A/B timing)
static bool virtual_pivot() {
  static const bool real = getenv("KISS_B200_REAL_PIVOT") != nullptr;
  return !real;
}

template <int NW, int R, bool WIDE>
static size_t med_smem_bytes() {
  const size_t slots = (size_t)(NW == 1 ? 8 : 1) * NW * R * 32;
  return slots * ((WIDE ? 16 : 8) + 12) + (NW > 1 ? MED_VCACHE * 8 + MED_NBK * 4 : 0);
}
template <int NW, int R, bool UNB, bool WIDE>
static void groups_medium_class(int pk, u32 cls, u32* pos, const u32* med, u32 medN, const u32* cg, const u32* ce,
                                SortParams sp, u64 h, cudaStream_t st) {
  const u32 ngroups = cg[cls];
  if (!ngroups) return;
  // (per launch: the attribute belongs to the current device, and a process may drive several)
  CK(cudaFuncSetAttribute(k_groups_medium<NW, R, UNB, WIDE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                          (int)med_smem_bytes<NW, R, WIDE>()));
  ProfScope ps(pk, st);
  if (ps.p) ps.p->bytes[pk] += 16ull * ce[cls];  // position read + write, one 8-byte key word
  u32 grid = NW == 1 ? (ngroups + 7) / 8 : ngroups;
  if (grid > 1048576u) grid = 1048576u;
  static const u32 rank_scale = [] { const char* e = getenv("KISS_B200_RANK_SCALE"); return e ? (u32)atoi(e) : 6u; }();  // C3: 2: 22.2 ms, 4: 20.5, 8: 19.8, 16: 21.7, 64: 41 (largest class)
  static const u32 bucket_cost = [] { const char* e = getenv("KISS_B200_BUCKET_COST"); return e ? (u32)atoi(e) : 128u; }();  // 0: the network in every first round
  static const u32 fin_max = [] { const char* e = getenv("KISS_B200_FIN_MAX"); return e ? (u32)atoi(e) : 4u; }();  // 0: off; C3 medium classes in all: 0: 27.1 ms, 4: 24.8, 8: 26.5, 16: 33.3
  k_groups_medium<NW, R, UNB, WIDE><<<grid, MED_THREADS(NW), med_smem_bytes<NW, R, WIDE>(), st>>>(
      pos, med + 2 * med_class_off(medN, cls), ngroups, sp, h, rank_scale, (NW > 1 && virtual_pivot()) ? 1u : 0u,
      bucket_cost, fin_max, g_med_dbg ? g_med_dbg + 64 * cls : nullptr);
  CK_LAUNCH(); g_launches++;
}
template <bool UNB, bool WIDE>
static void groups_medium_launch(u32* pos, const u32* med, u32 medN, const u32* cg, const u32* ce, SortParams sp, u64 h,
                                 cudaStream_t st) {
  groups_medium_class<1, 1, UNB, WIDE>(PK_GROUPS_MEDIUM, 0, pos, med, medN, cg, ce, sp, h, st);
  groups_medium_class<1, 2, UNB, WIDE>(PK_GROUPS_MEDIUM_64, 1, pos, med, medN, cg, ce, sp, h, st);
  groups_medium_class<1, 4, UNB, WIDE>(PK_GROUPS_MEDIUM_128, 2, pos, med, medN, cg, ce, sp, h, st);
  groups_medium_class<1, 8, UNB, WIDE>(PK_GROUPS_MEDIUM_256, 3, pos, med, medN, cg, ce, sp, h, st);
  groups_medium_class<8, 2, UNB, WIDE>(PK_GROUPS_MEDIUM_512, 4, pos, med, medN, cg, ce, sp, h, st);
  // 16 warps x 4 keys per thread (2 blocks = 32 warps per SM): 36.6 ms against 40.6 ms for 8 warps x 8 keys on C3
  groups_medium_class<16, 4, UNB, WIDE>(PK_GROUPS_MEDIUM_2048, 5, pos, med, medN, cg, ce, sp, h, st);
}

void groups_medium(u32* pos, const u32* med, u32 medN, const u32* class_groups, const u32* class_elems, SortParams sp,
                   u64 h, cudaStream_t st) {
  if (h >= sp.D) return;
  // two-word keys (46 own bases per round, KISS_B200_WIDE_KEYS=1) measured slower on C3: 18.9 ms against 15.0 ms in the
  // largest class (a third fewer element-rounds, but every step of sorting and ranking costs 1.6x)
  static const bool narrow = getenv("KISS_B200_WIDE_KEYS") == nullptr;
  // the 12-bit q field of the bounded key holds depths up to 4095
  if (sp.D == KISS_UNBOUNDED || sp.D > 4000) groups_medium_launch<true, false>(pos, med, medN, class_groups, class_elems, sp, h, st);
  else if (narrow) groups_medium_launch<false, false>(pos, med, medN, class_groups, class_elems, sp, h, st);
  else groups_medium_launch<false, true>(pos, med, medN, class_groups, class_elems, sp, h, st);
}

// ------------------------------------------------------------------------------------------- big groups
// Groups larger than a block's shared memory go through PIVOT ROUNDS in global memory: every element of a big
// group is compared with the group's first element from the group's known depth on (pivot_key above: first
// differing base q, side, next 14 bases), the composite key [group | side | q | bases] is sorted with the stable
// LSD radix sort, and the classes of equal keys are the groups of the next level (tiny / medium / big again).
// Elements equal to their pivot up to D get unique keys (their index in the group): they are final and must not
// group again.  Long common prefixes are skipped in one round; the first version advanced 16 bases per round
// (23 rounds at k = 256 on repeat-rich text).
__global__ void k_big_sizes(const u32* __restrict__ big, u32 nbig, u32* __restrict__ sizes) {
  u32 g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g < nbig) sizes[g] = big[3 * g + 1];
}
void big_sizes(const u32* big, u32 nbig, u32* sizes, cudaStream_t st) {
  if (!nbig) return;
  ProfScope ps(PK_GROUPS_OTHER, st);
  k_big_sizes<<<(nbig + 255) / 256, 256, 0, st>>>(big, nbig, sizes);
  CK_LAUNCH(); g_launches++;
}

// one thread per element of the work list; boff = exclusive scan of the group sizes
template <bool UNB>
__global__ void __launch_bounds__(256) k_big_fill(const u32* __restrict__ pos_prev, const u32* __restrict__ slot_prev,
                                                 const u32* __restrict__ big, const u32* __restrict__ boff, u32 nbig,
                                                 u32 total, SortParams sp, u64* __restrict__ wkey,
                                                 u32* __restrict__ wpos, u32* __restrict__ wslot, u32 vpivot) {
  constexpr int SIDE_SHIFT = UNB ? 35 : 40, G_SHIFT = UNB ? 37 : 42, DROP = UNB ? 16 : 11;
  const u32 t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  u32 lo = 0, hi = nbig;  // last group with boff <= t
  while (hi - lo > 1) {
    u32 mid = lo + (hi - lo) / 2;
    if (boff[mid] <= t) lo = mid; else hi = mid;
  }
  const u32 g = lo, start = big[3 * g], size = big[3 * g + 1], depth = big[3 * g + 2], j = t - boff[g];
  const u32 p = pos_prev[start + j];
  u64 f = (MK_EQUAL << SIDE_SHIFT) | j;  // the pivot itself, and everything equal to it: unique, in order
  if (UNB) {
    if (j) {
      const u64 key = pivot_key_u(sp, p, pos_prev[start], depth) >> DROP;
      if (((key >> SIDE_SHIFT) & 3ull) != MK_EQUAL) f = key;
    }
  } else if (vpivot || j) {
    // (one-word keys here: the 14 bases after q; the radix passes of the round grow with the key width)
    u64 lo;
    const u64 key = pivot_key_b(sp, p, pos_prev[start], pos_prev[start + (size >> 1)], pos_prev[start + size - 1],
                                vpivot != 0, depth, lo) >> DROP;
    if (((key >> SIDE_SHIFT) & 3ull) != MK_EQUAL) f = key;
  }
  wkey[t] = ((u64)g << G_SHIFT) | f;
  wpos[t] = p;
  wslot[t] = slot_prev ? slot_prev[start + j] : start + j;
}

void big_fill(const u32* pos_prev, const u32* slot_prev, const u32* big, const u32* boff, u32 nbig, u32 total,
              SortParams sp, bool unbounded_layout, u64* wkey, u32* wpos, u32* wslot, cudaStream_t st) {
  if (!nbig || !total) return;
  ProfScope ps(PK_GROUPS_OTHER, st);
  u32 grid = (total + 255) / 256;
  if (unbounded_layout)
    k_big_fill<true><<<grid, 256, 0, st>>>(pos_prev, slot_prev, big, boff, nbig, total, sp, wkey, wpos, wslot, 0u);
  else
    k_big_fill<false><<<grid, 256, 0, st>>>(pos_prev, slot_prev, big, boff, nbig, total, sp, wkey, wpos, wslot,
                                            virtual_pivot() ? 1u : 0u);
  CK_LAUNCH(); g_launches++;
}

__global__ void k_big_writeback(u32* __restrict__ pos_main, const u32* __restrict__ wslot,
                                const u32* __restrict__ wpos, u32 total) {
  u32 t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < total) pos_main[wslot[t]] = wpos[t];
}
void big_writeback(u32* pos_main, const u32* wslot, const u32* wpos, u32 total, cudaStream_t st) {
  if (!total) return;
  ProfScope ps(PK_GROUPS_OTHER, st);
  k_big_writeback<<<(total + 255) / 256, 256, 0, st>>>(pos_main, wslot, wpos, total);
  CK_LAUNCH(); g_launches++;
}

__global__ void k_gather_keys(const u64* __restrict__ T, const u32* __restrict__ pos, u32 N, u32* __restrict__ key) {
  u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < N) key[i] = (u32)(word_at(T, pos[i]) >> 32);
}
void gather_keys(const u64* T, const u32* pos, u32 N, u32* key, cudaStream_t st) {
  if (!N) return;
  k_gather_keys<<<(N + 255) / 256, 256, 0, st>>>(T, pos, N, key);
  CK_LAUNCH(); g_launches++;
}

// ------------------------------------------------------------------------------------------- sharding
// dest[i] = number of splitters <= key[i]  (equal keys always share a destination, so tie refinement and the
// position tie-break stay local to one GPU); counts[d] accumulates the destination sizes
__global__ void __launch_bounds__(256) k_dest_by_splitters(const u32* __restrict__ key, u32 N, const u32* __restrict__ spl,
                                                          u32 nspl, u32* __restrict__ dest, u32* __restrict__ counts) {
  __shared__ u32 sspl[64], scnt[65];
  if (threadIdx.x < 64) sspl[threadIdx.x] = threadIdx.x < nspl ? spl[threadIdx.x] : 0xFFFFFFFFu;
  if (threadIdx.x < 65) scnt[threadIdx.x] = 0;
  __syncthreads();
  u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < N) {
    u32 k = key[i], d = 0;
    for (u32 s = 0; s < nspl; s++) d += (sspl[s] <= k);
    dest[i] = d;
    atomicAdd(&scnt[d], 1u);
  }
  __syncthreads();
  if (threadIdx.x <= nspl && scnt[threadIdx.x]) atomicAdd(&counts[threadIdx.x], scnt[threadIdx.x]);
}
void dest_by_splitters(const u32* key, u32 N, const u32* spl, u32 nspl, u32* dest, u32* counts, cudaStream_t st) {
  if (!N) return;
  k_dest_by_splitters<<<(N + 255) / 256, 256, 0, st>>>(key, N, spl, nspl, dest, counts);
  CK_LAUNCH(); g_launches++;
}

// ------------------------------------------------------------------------------------------- short suffixes
// Literal restatement of the reference's comparator lambda (kiss1_core.hpp:94-135) on the packed text,
// preceded by the first-10-bases bucket order of the counting sort that runs before it (:41-83,
// structs.hpp:175-184; zero padding past the end).
__device__ bool ref_less(const SortParams& sp, u32 i, u32 j) {
  const u64 n = sp.n, k = sp.k;
  u32 pi = (u32)(word_at(sp.T, i) >> 44), pj = (u32)(word_at(sp.T, j) >> 44);
  if (pi != pj) return pi < pj;
  u64 sorted_len = 0, a = i, b = j;
  while (sorted_len <= k && a + 125 <= n && b + 125 <= n) {  // :96-98
#pragma unroll
    for (int w = 0; w < 4; w++) {                            // 125 bases = 3 words + 29 bases
      u64 x = word_at(sp.T, a + 32 * w), y = word_at(sp.T, b + 32 * w);
      if (w == 3) { x &= ~0ull << 6; y &= ~0ull << 6; }
      if (x != y) return x < y;                              // :104-114
    }
    sorted_len += 125; a += 125; b += 125;                   // :117-119
  }
  while (sorted_len + 1 <= k && a < n && b < n) {            // :121-130
    u32 ca = base_at(sp.T, a), cb = base_at(sp.T, b);
    if (ca != cb) return ca < cb;
    sorted_len++; a++; b++;
  }
  if (sorted_len >= k) return a < b;                         // :131-133
  return a == n;                                             // :134
}

__global__ void __launch_bounds__(128) k_short_rank(const u32* __restrict__ shorts, u32 ns, SortParams sp,
                                                   u32* __restrict__ rank0) {
  u32 s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= ns) return;
  u32 me = shorts[s], r = 0;
  for (u32 t = 0; t < ns; t++)
    if (t != s && ref_less(sp, shorts[t], me)) r++;
  rank0[s] = r;
}
// make the ranks a permutation even if the comparator is not a strict weak order
__global__ void __launch_bounds__(128) k_short_place(const u32* __restrict__ shorts, u32 ns, const u32* __restrict__ rank0,
                                                    u32* __restrict__ sorted_shorts) {
  u32 s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= ns) return;
  u32 r0 = rank0[s], r = 0;
  for (u32 t = 0; t < ns; t++) {
    u32 rt = rank0[t];
    if (rt < r0 || (rt == r0 && t < s)) r++;
  }
  sorted_shorts[r] = shorts[s];
}

// Many short suffixes (a bounded k that is a large part of n): a bitonic network in global memory with the same
// comparator — O(ns log^2 ns) comparisons where the all-pairs ranking above is O(ns^2).  The comparator is a total order
// on distinct positions whenever it is transitive; where it is not the reference's own std::sort result
// is unspecified too.  a[] is padded to a power of two with KISS_EMPTY = +infinity.
__global__ void __launch_bounds__(256) k_short_bitonic(u32* __restrict__ a, u32 npad, u32 k2, u32 j, SortParams sp) {
  const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npad) return;
  const u32 p = i ^ j;
  if (p <= i) return;
  const u32 x = a[i], y = a[p];
  const bool y_first = x == KISS_EMPTY ? y != KISS_EMPTY : (y != KISS_EMPTY && ref_less(sp, y, x));
  const bool asc = (i & k2) == 0;
  if (y_first == asc) { a[i] = y; a[p] = x; }
}
#define SHORT_ALLPAIRS_MAX 4096u

void shorts_order(const u32* shorts, u32 ns, SortParams sp, u32* rank0, u32* sorted_shorts, cudaStream_t st) {
  if (!ns) return;
  ProfScope ps(PK_SHORTS, st);
  if (ns > SHORT_ALLPAIRS_MAX) {
    // rank0 is the network's work array here: shorts_scratch_len(ns) entries
    u32 npad = 1;
    while (npad < ns) npad <<= 1;
    CK(cudaMemsetAsync(rank0, 0xFF, (u64)npad * sizeof(u32), st));
    CK(cudaMemcpyAsync(rank0, shorts, (u64)ns * sizeof(u32), cudaMemcpyDeviceToDevice, st));
    for (u32 k2 = 2; k2 <= npad; k2 <<= 1)
      for (u32 j = k2 >> 1; j > 0; j >>= 1) {
        k_short_bitonic<<<(npad + 255) / 256, 256, 0, st>>>(rank0, npad, k2, j, sp);
        CK_LAUNCH();
        g_launches++;
      }
    CK(cudaMemcpyAsync(sorted_shorts, rank0, (u64)ns * sizeof(u32), cudaMemcpyDeviceToDevice, st));
    return;
  }
  k_short_rank<<<(ns + 127) / 128, 128, 0, st>>>(shorts, ns, sp, rank0);
  CK_LAUNCH();
  k_short_place<<<(ns + 127) / 128, 128, 0, st>>>(shorts, ns, rank0, sorted_shorts);
  CK_LAUNCH();
  g_launches += 2;
}

// ins[s] = first index of the sorted long list whose suffix is not before short s
__global__ void __launch_bounds__(128) k_short_ins(const u32* __restrict__ sorted_shorts, u32 ns,
                                                  const u32* __restrict__ longs, u32 nl, SortParams sp,
                                                  u32* __restrict__ ins) {
  u32 s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= ns) return;
  u32 me = sorted_shorts[s], lo = 0, hi = nl;
  while (lo < hi) {
    u32 mid = lo + (hi - lo) / 2;
    if (ref_less(sp, longs[mid], me)) lo = mid + 1; else hi = mid;
  }
  ins[s] = lo;
}
// running maximum: short s is emitted no earlier than the shorts before it
__global__ void k_short_mono(u32* __restrict__ ins, u32 ns) {
  u32 mx = 0;
  for (u32 s = 0; s < ns; s++) {
    mx = max(mx, ins[s]);
    ins[s] = mx;
  }
}

void shorts_insertions(const u32* sorted_shorts, u32 ns, const u32* longs, u32 nl, SortParams sp, u32* ins,
                       cudaStream_t st) {
  if (!ns) return;
  ProfScope ps(PK_SHORTS, st);
  k_short_ins<<<(ns + 127) / 128, 128, 0, st>>>(sorted_shorts, ns, longs, nl, sp, ins);
  CK_LAUNCH();
  k_short_mono<<<1, 1, 0, st>>>(ins, ns);
  CK_LAUNCH();
  g_launches += 2;
}

}  // namespace kiss
