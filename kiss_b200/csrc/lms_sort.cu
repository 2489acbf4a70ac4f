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
//   tiny   (<= 8)     one thread, insertion sort with the word-wise suffix comparator
//   medium (<= 2048)  one block, bitonic sort in shared memory with the same comparator
//   big               radix rounds: composite key (group id, next 16 bases) sorted with the LSD sort, the
//                     resulting sub-groups go through the same three cases at the next depth
// All paths are stable with respect to text position, which is the reference's final tie-break (:131-133).
#include "common.cuh"
#include "kernels.h"

namespace kiss {

// true iff suffix p sorts before suffix q; both are known equal on their first h bases.
// Bounded mode: both suffixes have at least D bases.  Unbounded (D = KISS_UNBOUNDED): the shorter suffix
// first when one is a prefix of the other (kiss1_core.hpp:134).
__device__ __forceinline__ bool suffix_less(const SortParams& sp, u32 p, u32 q, u64 h) {
  for (u64 off = h; off < sp.D; off += 32) {
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
// One block looks at 256 consecutive sorted keys (+ a halo).  Every element whose key equals a neighbour's
// fetches its next 32 bases in parallel (independent gathers, no dependent chains); then the first element of
// each group of <= TINY_MAX orders the group in shared memory by (next word, deeper words, position).
#define TG_THREADS 256
#define TG_ITEMS 8
#define TG_TILE (TG_THREADS * TG_ITEMS)
template <typename K>
__global__ void __launch_bounds__(TG_THREADS) k_groups_tiny(const K* __restrict__ keys, u32* __restrict__ pos, u32 N,
                                                           SortParams sp, u64 h, GroupLists gl) {
  __shared__ K skey[TG_TILE + TINY_MAX + 2];   // keys[base-1 .. base+TG_TILE+TINY_MAX]
  __shared__ u32 spos[TG_TILE + TINY_MAX];
  __shared__ u64 sw[TG_TILE + TINY_MAX];
  __shared__ unsigned short sheads[TG_TILE / 2];  // tile-local index of every group head (a group has >= 2 elements)
  __shared__ u32 s_tied, s_nheads;
  const u32 tid = threadIdx.x;
  if (tid == 0) { s_tied = 0; s_nheads = 0; }
  const u64 base = (u64)blockIdx.x * TG_TILE;
  for (u32 j = tid; j < TG_TILE + TINY_MAX + 2; j += TG_THREADS) {
    u64 idx = base + j;  // = element index + 1
    skey[j] = (idx >= 1 && idx - 1 < N) ? keys[idx - 1] : (K)0;
  }
  __syncthreads();
  const bool deep = h < sp.D;
  const u64 rem = sp.D - h;
  {
    // all tied elements of the tile fetch position and next word; loads of one thread are independent.
    // Group heads (tied with the next element, not with the previous one) are queued for the second phase.
    u32 p[TG_ITEMS + 1];
    bool tied[TG_ITEMS + 1];
#pragma unroll
    for (int it = 0; it <= TG_ITEMS; it++) {
      u32 t = tid + it * TG_THREADS;
      u64 e = base + t;
      tied[it] = false;
      p[it] = 0;
      bool head = false;
      if (t < TG_TILE + TINY_MAX && e < N) {
        K k = skey[t + 1];
        bool prev = e > 0 && skey[t] == k, next = e + 1 < N && skey[t + 2] == k;
        tied[it] = prev || next;
        head = next && !prev && t < TG_TILE;
        if (tied[it] && deep) p[it] = pos[e];
      }
      u32 hb = __ballot_sync(0xffffffffu, head);
      if (hb) {
        u32 lane = tid & 31u, slot = 0;
        if (lane == 0) slot = atomicAdd(&s_nheads, (u32)__popc(hb));
        slot = __shfl_sync(0xffffffffu, slot, 0);
        if (head) sheads[slot + __popc(hb & ((1u << lane) - 1u))] = (unsigned short)t;
      }
    }
    if (deep) {
#pragma unroll
      for (int it = 0; it <= TG_ITEMS; it++) {
        if (tied[it]) {
          u32 t = tid + it * TG_THREADS;
          u64 wv = word_at(sp.T, (u64)p[it] + h);
          if (rem < 32) wv &= ~0ull << (64 - 2 * rem);
          spos[t] = p[it];
          sw[t] = wv;
        }
      }
    }
  }
  __syncthreads();
  const u32 nheads = s_nheads;
  for (u32 hi = tid; hi < nheads; hi += TG_THREADS) {
    const u32 t = sheads[hi];
    const u64 e = base + t;
    const K k = skey[t + 1];
    u32 g = 2;
    while (g <= TINY_MAX && e + g < N && skey[t + 1 + g] == k) g++;
    if (g > TINY_MAX) {
      u32 idx = atomicAdd(&gl.counters[0], 1u);
      if (idx < gl.gcap) gl.glist[idx] = (u32)e; else gl.counters[3] = 1;
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

template <typename K>
__global__ void __launch_bounds__(128) k_groups_classify(const K* __restrict__ keys, u32 N, u32 ng, GroupLists gl) {
  u32 g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= ng) return;
  u64 start = gl.glist[g];
  K k = keys[start];
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
  u32 size = (u32)(hi - start);
  atomicAdd(gl.tied, (u64)size);
  if (size <= MED_MAX) {
    u32 j = atomicAdd(&gl.counters[1], 1u);
    gl.med[2 * j] = (u32)start;
    gl.med[2 * j + 1] = size;
  } else {
    u32 j = atomicAdd(&gl.counters[2], 1u);
    gl.big[2 * j] = (u32)start;
    gl.big[2 * j + 1] = size;
  }
}

void groups_tiny_u32(const u32* keys, u32* pos, u32 N, SortParams sp, u64 h, GroupLists gl, cudaStream_t st) {
  if (!N) return;
  ProfScope ps(PK_GROUPS_TINY, st);
  k_groups_tiny<u32><<<(N + TG_TILE - 1) / TG_TILE, TG_THREADS, 0, st>>>(keys, pos, N, sp, h, gl);
  CK_LAUNCH(); g_launches++;
}
void groups_tiny_u64(const u64* keys, u32* pos, u32 N, SortParams sp, u64 h, GroupLists gl, cudaStream_t st) {
  if (!N) return;
  ProfScope ps(PK_GROUPS_TINY, st);
  k_groups_tiny<u64><<<(N + TG_TILE - 1) / TG_TILE, TG_THREADS, 0, st>>>(keys, pos, N, sp, h, gl);
  CK_LAUNCH(); g_launches++;
}
void groups_classify_u32(const u32* keys, u32 N, u32 ng, GroupLists gl, cudaStream_t st) {
  if (!ng) return;
  ProfScope ps(PK_GROUPS_OTHER, st);
  k_groups_classify<u32><<<(ng + 127) / 128, 128, 0, st>>>(keys, N, ng, gl);
  CK_LAUNCH(); g_launches++;
}
void groups_classify_u64(const u64* keys, u32 N, u32 ng, GroupLists gl, cudaStream_t st) {
  if (!ng) return;
  ProfScope ps(PK_GROUPS_OTHER, st);
  k_groups_classify<u64><<<(ng + 127) / 128, 128, 0, st>>>(keys, N, ng, gl);
  CK_LAUNCH(); g_launches++;
}

// ------------------------------------------------------------------------------------------- medium groups
// One team (a warp for groups <= 64, a block above) refines one group in shared memory by rounds of 32 bases:
// every element that still shares its segment with another fetches its next word (independent gathers), the
// group is bitonic-sorted on (segment, word, position) and the segments are split where the words differ.
// A round whose words are all equal inside their segments skips the sort (long shared prefixes — copies of a
// repeat — cost one gather per round and nothing else), and the loop ends as soon as every segment is a
// singleton.  Gathers per element: at most (D - h) / 32, instead of one dependent chain per comparison.
#define MEDW_MAX 64u
struct MedView {
  u64* key;
  u32* pos;
  u32* seg;
};

template <u32 GT>
__device__ __forceinline__ void team_sync() {
  if (GT == 32) __syncwarp(); else __syncthreads();
}
template <u32 GT>
__device__ __forceinline__ bool team_or(bool v) {
  if (GT == 32) return __any_sync(0xffffffffu, v);
  return __syncthreads_or(v);
}

// element i before element j at depth `off` (both in the same array; padding has seg = 0xFFFFFFFF)
__device__ __forceinline__ bool med_less(const MedView& m, u32 i, u32 j, u64 off, u64 n) {
  u32 si = m.seg[i], sj = m.seg[j];
  if (si != sj) return si < sj;
  u32 pi = m.pos[i], pj = m.pos[j];
  bool ei = (u64)pi + off >= n, ej = (u64)pj + off >= n;  // suffix ended: a prefix of the others, hence first
  if (ei | ej) return (ei && ej) ? pi > pj : ei;
  u64 ki = m.key[i], kj = m.key[j];
  if (ki != kj) return ki < kj;
  return pi < pj;
}

template <u32 GT>
__device__ __forceinline__ void med_refine(const MedView& m, u32 t, u32 size, u32 P, const SortParams& sp, u64 h,
                                           u32* s_scan) {
  for (u64 off = h; off < sp.D; off += 32) {
    const u64 rem = sp.D - off;
    const u64 mk = rem < 32 ? ~0ull << (64 - 2 * rem) : ~0ull;
    bool anyact = false;
    for (u32 i = t; i < size; i += GT) {
      u32 sg = m.seg[i];
      bool act = sg != i || (i + 1 < size && m.seg[i + 1] == sg);
      u64 key = 0;
      if (act) {
        anyact = true;
        u64 p = m.pos[i];
        if (p + off < sp.n) key = word_at(sp.T, p + off) & mk;
      }
      m.key[i] = key;
    }
    if (!team_or<GT>(anyact)) break;
    bool need = false;
    for (u32 i = t; i < size; i += GT)
      if (m.seg[i] != i) {  // not a segment head: same segment as i - 1
        bool e = (u64)m.pos[i] + off >= sp.n || (u64)m.pos[i - 1] + off >= sp.n;
        need |= e || m.key[i] != m.key[i - 1];
      }
    if (!team_or<GT>(need)) continue;
    for (u32 k2 = 2; k2 <= P; k2 <<= 1) {
      for (u32 j = k2 >> 1; j > 0; j >>= 1) {
        for (u32 i = t; i < P; i += GT) {
          u32 x = i ^ j;
          if (x > i) {
            bool asc = (i & k2) == 0;
            if (med_less(m, x, i, off, sp.n) == asc) {
              u64 kk = m.key[i]; m.key[i] = m.key[x]; m.key[x] = kk;
              u32 pp = m.pos[i]; m.pos[i] = m.pos[x]; m.pos[x] = pp;
              u32 ss = m.seg[i]; m.seg[i] = m.seg[x]; m.seg[x] = ss;
            }
          }
        }
        team_sync<GT>();
      }
    }
    // split the segments: thread t owns E consecutive slots; head index = running maximum
    const u32 E = P > GT ? P / GT : 1u;
    const u32 i0 = t * E;
    u32 heads = 0;
    for (u32 e = 0; e < E; e++) {
      u32 i = i0 + e;
      if (i < size) {
        bool hd = i == 0 || m.seg[i] != m.seg[i - 1] || m.key[i] != m.key[i - 1] ||
                  (u64)m.pos[i] + off >= sp.n || (u64)m.pos[i - 1] + off >= sp.n;
        heads |= (hd ? 1u : 0u) << e;
      }
    }
    u32 v = heads ? i0 + (31u - (u32)__clz((int)heads)) + 1u : 0u;  // last head in my slots, +1
    u32 incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      u32 u = __shfl_up_sync(0xffffffffu, incl, o);
      if ((int)(t & 31u) >= o) incl = max(incl, u);
    }
    u32 excl = __shfl_up_sync(0xffffffffu, incl, 1);
    if ((t & 31u) == 0) excl = 0;
    if (GT > 32) {
      if ((t & 31u) == 31u) s_scan[t >> 5] = incl;
      __syncthreads();
      for (u32 w = 0; w < (t >> 5); w++) excl = max(excl, s_scan[w]);
    } else {
      __syncwarp();
    }
    u32 cur = excl;  // head index + 1 of the segment running into my first slot
    for (u32 e = 0; e < E; e++) {
      u32 i = i0 + e;
      if (i < size) {
        if ((heads >> e) & 1u) cur = i + 1;
        m.seg[i] = cur - 1;
      }
    }
    team_sync<GT>();
  }
}

// GT = 32: one warp per group of <= MEDW_MAX; GT = 256: one block per larger group
template <u32 GT>
__global__ void __launch_bounds__(256) k_groups_medium(u32* __restrict__ pos, const u32* __restrict__ med, u32 nmed,
                                                      SortParams sp, u64 h) {
  constexpr u32 CAP = GT == 32 ? MEDW_MAX : MED_MAX;
  constexpr u32 TEAMS = 256 / GT;
  __shared__ u64 s_key[TEAMS * CAP];
  __shared__ u32 s_pos[TEAMS * CAP];
  __shared__ u32 s_seg[TEAMS * CAP];
  __shared__ u32 s_scan[8];
  const u32 team = threadIdx.x / GT, t = threadIdx.x % GT;
  MedView m{s_key + team * CAP, s_pos + team * CAP, s_seg + team * CAP};
  for (u32 gi = blockIdx.x * TEAMS + team; gi < nmed; gi += gridDim.x * TEAMS) {
    const u32 start = med[2 * gi], size = med[2 * gi + 1];
    const bool mine = GT == 32 ? size <= MEDW_MAX : size > MEDW_MAX;
    if (!mine) continue;  // (uniform per team)
    u32 P = 2;
    while (P < size) P <<= 1;
    for (u32 i = t; i < P; i += GT) {
      m.pos[i] = i < size ? pos[start + i] : KISS_EMPTY;
      m.seg[i] = i < size ? 0u : 0xFFFFFFFFu;
    }
    team_sync<GT>();
    med_refine<GT>(m, t, size, P, sp, h, s_scan);
    team_sync<GT>();
    for (u32 i = t; i < size; i += GT) pos[start + i] = m.pos[i];
    team_sync<GT>();
  }
}

void groups_medium(u32* pos, const u32* med, u32 nmed, SortParams sp, u64 h, cudaStream_t st) {
  if (!nmed || h >= sp.D) return;
  ProfScope ps(PK_GROUPS_MEDIUM, st);
  u32 gw = (nmed + 7) / 8, gb = nmed;
  if (gw > 1048576u) gw = 1048576u;
  if (gb > 1048576u) gb = 1048576u;
  k_groups_medium<32><<<gw, 256, 0, st>>>(pos, med, nmed, sp, h);
  CK_LAUNCH();
  k_groups_medium<256><<<gb, 256, 0, st>>>(pos, med, nmed, sp, h);
  CK_LAUNCH();
  g_launches += 2;
}

// ------------------------------------------------------------------------------------------- big groups
__global__ void k_big_sizes(const u32* __restrict__ big, u32 nbig, u32* __restrict__ sizes) {
  u32 g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g < nbig) sizes[g] = big[2 * g + 1];
}
void big_sizes(const u32* big, u32 nbig, u32* sizes, cudaStream_t st) {
  if (!nbig) return;
  ProfScope ps(PK_GROUPS_OTHER, st);
  k_big_sizes<<<(nbig + 255) / 256, 256, 0, st>>>(big, nbig, sizes);
  CK_LAUNCH(); g_launches++;
}

__global__ void __launch_bounds__(256) k_big_fill(const u32* __restrict__ pos_prev, const u32* __restrict__ slot_prev,
                                                 const u32* __restrict__ big, const u32* __restrict__ boff, u32 nbig,
                                                 SortParams sp, u64 h, u64* __restrict__ wkey, u32* __restrict__ wpos,
                                                 u32* __restrict__ wslot) {
  for (u32 g = blockIdx.x; g < nbig; g += gridDim.x) {
    u32 start = big[2 * g], size = big[2 * g + 1], off = boff[g];
    for (u32 j = threadIdx.x; j < size; j += blockDim.x) {
      u32 p = pos_prev[start + j];
      u64 c33;
      if ((u64)p + h >= sp.n) {
        c33 = 0;  // the suffix ended: it is a proper prefix of the others in its group, hence first
      } else {
        u32 c = (u32)(word_at(sp.T, (u64)p + h) >> 32);
        u64 rem = sp.D - h;
        if (rem < 16) c &= ~0u << (32 - 2 * (u32)rem);
        c33 = 1ull + c;
      }
      wkey[off + j] = ((u64)g << 33) | c33;
      wpos[off + j] = p;
      wslot[off + j] = slot_prev ? slot_prev[start + j] : start + j;
    }
  }
}

void big_fill(const u32* pos_prev, const u32* slot_prev, const u32* big, const u32* boff, u32 nbig, SortParams sp,
              u64 h, u64* wkey, u32* wpos, u32* wslot, cudaStream_t st) {
  if (!nbig) return;
  ProfScope ps(PK_GROUPS_OTHER, st);
  u32 grid = nbig < 65535u * 16u ? nbig : 65535u * 16u;
  k_big_fill<<<grid, 256, 0, st>>>(pos_prev, slot_prev, big, boff, nbig, sp, h, wkey, wpos, wslot);
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

void shorts_order(const u32* shorts, u32 ns, SortParams sp, u32* rank0, u32* sorted_shorts, cudaStream_t st) {
  if (!ns) return;
  ProfScope ps(PK_SHORTS, st);
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
