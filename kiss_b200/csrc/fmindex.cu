// fmindex.cu — the step after the path: BWT, occurrence tables and sampled suffix array from
// the suffix array, on the device, so that the 4(n+1)-byte array never crosses PCIe (1.75 bytes per suffix do instead).
//
// Replaces FMIndex::build_occ / build_bwt / build_sa (include/biovoltron/algo/align/exact_match/fm_index.hpp:278-307,
// 311-328, 332-372) as instantiated by the reference's fmindex_build command (include/command/fmindex_build.hpp:27-34:
// FMIndex<SA_INTV = 4, uint32_t, KISS1Sorter<uint32_t>>{.LOOKUP_LEN = 0}).  With N = n + 1 entries:
//   bwt_    DibitVector: entry i = S[SA[i] - 1] (0 where SA[i] == 0, whose index is pri_), element e in byte e / 4 at
//           bits 2 (e % 4)  (xbit_vector.hpp:21-27)
//   occ_    first: per 256 entries the exclusive prefix of the symbol counts (u32 x 4, N / 256 + 1 rows);
//           second: per 16 entries the counts since the row's 256-boundary (u8 x 4, N / 16 + 1 rows); the entry with
//           SA == 0 counts for nothing (:295-297); cnt_[c] = 1 + number of symbols smaller than c
//   b_      one bit per entry: SA[i] % SA_INTV == 0; b_occ_: per 64 entries the exclusive prefix of its popcounts
//   sa_     the entries with SA[i] % SA_INTV == 0, in suffix-array order
// One pass over SA (k_fm_scan: one block = one 256-entry row; the base before a suffix comes from the context word the
// induction left beside every entry, so only one entry in sixteen gathers from the text: C3 69.6 -> see DESIGN.md),
// three small scans, one more pass over SA for the compaction (k_fm_sample).  Algorithmic bytes: 2 x 4N (SA
// reads) + N / 4 (gathered text, ideal) + 1.75 N written.
#include <cstdlib>
#include "common.cuh"
#include "kernels.h"

namespace kiss {

#define FM_ROW 256u

__global__ void __launch_bounds__(FM_ROW) k_fm_scan(const u32* __restrict__ SA, const u32* __restrict__ CTX, u64 N,
                                                   const u64* __restrict__ T, u32 intv,
                                                   u32* __restrict__ bwt32, u32* __restrict__ tot /* [4][n1: the plane stride] */, u64 n1,
                                                   u8* __restrict__ occ2, u64 n2, u32* __restrict__ b32,
                                                   u32* __restrict__ bocc_raw, u32* __restrict__ pri) {
  __shared__ u32 s_seg[16][4];  // counts per 16-entry segment and symbol
  __shared__ u32 s_hit[8];      // sampled entries per warp
  const u32 t = threadIdx.x, lane = t & 31u, w = t >> 5;
  const u64 i = (u64)blockIdx.x * FM_ROW + t;
  const bool valid = i < N;
  const u32 v = valid ? SA[i] : 1u;
  const bool counted = valid && v != 0;
  // the base before the suffix: the induction left it in the low bits of the entry's context word (induce.cu); the text
  // is gathered only where a context had run out (about one entry in sixteen)
  u32 c = 0;
  if (counted) {
    const u32 cx = CTX ? CTX[i] : 1u;
    c = cx > 1u ? (cx & 3u) : base_at(T, (u64)v - 1);
  }
  if (valid && v == 0) *pri = (u32)i;
  // bwt: 4 entries per byte, 16 per word; lanes 0 and 16 of a warp store one word each
  u32 x = c << (2 * (lane & 15u));
  x |= __shfl_xor_sync(0xffffffffu, x, 1);
  x |= __shfl_xor_sync(0xffffffffu, x, 2);
  x |= __shfl_xor_sync(0xffffffffu, x, 4);
  x |= __shfl_xor_sync(0xffffffffu, x, 8);
  if ((lane & 15u) == 0 && (i & ~15ull) < N) bwt32[i >> 4] = x;
  // symbol counts per half warp = per 16-entry segment
#pragma unroll
  for (u32 s = 0; s < 4; s++) {
    const u32 m = __ballot_sync(0xffffffffu, counted && c == s);
    if (lane == 0) {
      s_seg[2 * w][s] = (u32)__popc(m & 0xFFFFu);
      s_seg[2 * w + 1][s] = (u32)__popc(m >> 16);
    }
  }
  // sampled-entry bits; b_occ raw = popcount per 64 entries = per warp pair
  if (b32) {
    const u32 hit = __ballot_sync(0xffffffffu, valid && v % intv == 0);
    if (lane == 0) {
      if ((i & ~31ull) < N) b32[i >> 5] = hit;
      s_hit[w] = (u32)__popc(hit);
    }
  }
  __syncthreads();
  if (b32 && t < 4) {
    const u64 j = (u64)blockIdx.x * 4 + t;  // 64-entry block
    if (j * 64 < N) bocc_raw[j] = s_hit[2 * t] + s_hit[2 * t + 1];
  }
  if (t < 4) {
    u32 run = 0;
    for (u32 s = 0; s < 16; s++) {
      const u64 j = (u64)blockIdx.x * 16 + s;
      const u32 cur = s_seg[s][t];
      if (j < n2) occ2[4 * j + t] = (u8)run;
      run += cur;
    }
    tot[(u64)t * n1 + blockIdx.x] = run;
  }
}

// occ1 rows from the four scanned planes
__global__ void k_fm_occ1(const u32* __restrict__ tot, u64 stride, u64 n1, u32* __restrict__ occ1) {
  const u64 b = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n1) return;
  *reinterpret_cast<uint4*>(occ1 + 4 * b) = make_uint4(tot[b], tot[stride + b], tot[2 * stride + b], tot[3 * stride + b]);
}

// sa_[b_occ[i / 64] + (sampled entries of the 64-block before i)] = SA[i] for the sampled entries
__global__ void __launch_bounds__(256) k_fm_sample(const u32* __restrict__ SA, u64 N, u32 intv, const u32* __restrict__ b32,
                                                  const u32* __restrict__ bocc, u32* __restrict__ sa_out) {
  const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const u32 v = SA[i];
  if (v % intv != 0) return;
  const u64 w0 = (i >> 6) * 2;  // first 32-bit word of the 64-entry block
  const u32 bit = (u32)(i & 63);
  u32 before = bit >= 32 ? (u32)__popc(b32[w0]) + (u32)__popc(b32[w0 + 1] & ((1u << (bit - 32)) - 1u))
                         : (u32)__popc(b32[w0] & ((1u << bit) - 1u));
  sa_out[bocc[i >> 6] + before] = v;
}

// ---- the same scan with FOUR consecutive entries per thread (two 16-byte loads): the one-entry kernel above keeps 2048
// four-byte loads in flight per SM, half of what the HBM latency needs, and ran at 1 TB/s (profiles/r2_fm_kernels.txt).
// One block = 1024 entries = four occ1 rows; a row is two warps, a 16-entry occ2 segment four lanes, a b_ word eight,
// a b_occ block sixteen.  Needs SA (and CTX) 16-byte aligned; the launcher falls back to the kernel above otherwise.
// v % intv == 0 (a mask when the interval is a power of two, as the reference's SA_INTV = 4 is: the division was a quarter
// of the scan's instructions, profiles/r2_ncu_fm_scan4.txt)
__device__ __forceinline__ bool fm_sampled(u32 v, u32 intv, u32 mask) { return mask != 0xFFFFFFFFu ? (v & mask) == 0u : v % intv == 0u; }
#define FM4_THREADS 256u
#define FM4_TILE (4u * FM4_THREADS)
__global__ void __launch_bounds__(FM4_THREADS) k_fm_scan4(const u32* __restrict__ SA, const u32* __restrict__ CTX, u64 N,
                                                         const u64* __restrict__ T, u32 intv, u32* __restrict__ bwt32,
                                                         u32* __restrict__ tot, u64 n1p, u64 n1, u32* __restrict__ occ2w, u64 n2,
                                                         u32* __restrict__ b32, u32* __restrict__ bocc_raw, u32* __restrict__ pri) {
  __shared__ u64 s_wtot[FM4_THREADS / 32];  // symbol counts per warp, 4 x 16 bits
  const u32 t = threadIdx.x, lane = t & 31u, w = t >> 5;
  const u64 i0 = (u64)blockIdx.x * FM4_TILE + 4ull * t;
  u32 v[4] = {1u, 1u, 1u, 1u}, cx[4] = {1u, 1u, 1u, 1u};
  if (i0 + 4 <= N) {
    const uint4 q = *reinterpret_cast<const uint4*>(SA + i0);
    v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
    if (CTX) {
      const uint4 c = *reinterpret_cast<const uint4*>(CTX + i0);
      cx[0] = c.x; cx[1] = c.y; cx[2] = c.z; cx[3] = c.w;
    }
  } else {
#pragma unroll
    for (int e = 0; e < 4; e++)
      if (i0 + e < N) { v[e] = SA[i0 + e]; if (CTX) cx[e] = CTX[i0 + e]; }
  }
  const u32 imask = (intv & (intv - 1u)) == 0u ? intv - 1u : 0xFFFFFFFFu;
  u32 byte = 0, hit4 = 0;
  u64 pc = 0;
#pragma unroll
  for (int e = 0; e < 4; e++) {
    const bool valid = i0 + e < N, counted = valid && v[e] != 0;
    u32 c = 0;
    if (counted) c = cx[e] > 1u ? (cx[e] & 3u) : base_at(T, (u64)v[e] - 1);
    if (valid && v[e] == 0) *pri = (u32)(i0 + e);
    byte |= c << (2 * e);
    if (counted) pc += 1ull << (16 * c);
    if (valid && fm_sampled(v[e], intv, imask)) hit4 |= 1u << e;
  }
  // bwt: 16 entries per word = four lanes
  u32 x = byte << (8 * (lane & 3u));
  x |= __shfl_xor_sync(0xffffffffu, x, 1);
  x |= __shfl_xor_sync(0xffffffffu, x, 2);
  if ((lane & 3u) == 0 && i0 < N) bwt32[i0 >> 4] = x;
  // symbol counts: inclusive scan over the lanes, warp totals to shared memory
  u64 incl = pc;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const u64 y = __shfl_up_sync(0xffffffffu, incl, o);
    if ((int)lane >= o) incl += y;
  }
  if (lane == 31) s_wtot[w] = incl;
  // sampled-entry bits: a b_ word = eight lanes, a b_occ block = sixteen
  if (b32) {
    u32 hb = hit4 << (4 * (lane & 7u));
    hb |= __shfl_xor_sync(0xffffffffu, hb, 1);
    hb |= __shfl_xor_sync(0xffffffffu, hb, 2);
    hb |= __shfl_xor_sync(0xffffffffu, hb, 4);
    if ((lane & 7u) == 0 && i0 < N) b32[i0 >> 5] = hb;
    u32 hc = (u32)__popc(hit4);
    hc += __shfl_xor_sync(0xffffffffu, hc, 1);
    hc += __shfl_xor_sync(0xffffffffu, hc, 2);
    hc += __shfl_xor_sync(0xffffffffu, hc, 4);
    hc += __shfl_xor_sync(0xffffffffu, hc, 8);
    if ((lane & 15u) == 0 && i0 < N) bocc_raw[i0 >> 6] = hc;
  }
  __syncthreads();
  // occ2: counts since the row's start before every 16-entry segment (its first lane writes the four bytes at once)
  const u64 before_seg = __shfl_sync(0xffffffffu, incl, (lane & ~3u) - 1u & 31u);  // inclusive count of the lane before the segment
  if ((lane & 3u) == 0) {
    u64 p = lane ? before_seg : 0ull;
    if (w & 1u) p += s_wtot[w - 1];  // the second warp of the row
    const u64 j = i0 >> 4;
    if (j < n2)
      occ2w[j] = (u32)(p & 0xFFu) | ((u32)((p >> 16) & 0xFFu) << 8) | ((u32)((p >> 32) & 0xFFu) << 16) | ((u32)((p >> 48) & 0xFFu) << 24);
  }
  if (t < 16) {
    const u32 row = t >> 2, sym = t & 3u;
    const u64 r = (u64)blockIdx.x * 4 + row;
    if (r < n1 && r * FM_ROW < N) {
      const u64 sum = s_wtot[2 * row] + s_wtot[2 * row + 1];
      tot[(u64)sym * n1p + r] = (u32)((sum >> (16 * sym)) & 0xFFFFu);
    }
  }
}

// the compaction with four entries per thread
__global__ void __launch_bounds__(256) k_fm_sample4(const u32* __restrict__ SA, u64 N, u32 intv, const u32* __restrict__ b32,
                                                   const u32* __restrict__ bocc, u32* __restrict__ sa_out) {
  const u64 i0 = ((u64)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i0 >= N) return;
  u32 v[4] = {1u, 1u, 1u, 1u};
  if (i0 + 4 <= N) {
    const uint4 q = *reinterpret_cast<const uint4*>(SA + i0);
    v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
  } else {
    for (int e = 0; e < 4; e++)
      if (i0 + e < N) v[e] = SA[i0 + e];
  }
  const u32 imask = (intv & (intv - 1u)) == 0u ? intv - 1u : 0xFFFFFFFFu;
  bool any = false;
#pragma unroll
  for (int e = 0; e < 4; e++) any |= (i0 + e < N) && fm_sampled(v[e], intv, imask);
  if (!any) return;
  const u64 w0 = (i0 >> 6) * 2;
  const u64 bits = (u64)b32[w0] | ((u64)b32[w0 + 1] << 32);
  const u32 base = bocc[i0 >> 6];
#pragma unroll
  for (int e = 0; e < 4; e++) {
    if (i0 + e < N && fm_sampled(v[e], intv, imask)) {
      const u32 bit = (u32)((i0 + e) & 63);
      sa_out[base + (u32)__popcll(bits & ((1ull << bit) - 1ull))] = v[e];
    }
  }
}

void fm_scan(const u32* SA, const u32* CTX, u64 N, const u64* T, u32 intv, u32* bwt32, u32* tot, u64 n1, u8* occ2, u64 n2, u32* b32,
             u32* bocc_raw, u32* pri, cudaStream_t st) {
  static const bool one = getenv("KISS_B200_FM_ONE") != nullptr;  // A/B: one entry per thread
  const bool aligned = (((uintptr_t)SA | (uintptr_t)CTX | (uintptr_t)occ2) & 15) == 0;
  if (aligned && !one) {
    const u64 blocks = (N + FM4_TILE - 1) / FM4_TILE;
    k_fm_scan4<<<(unsigned)blocks, FM4_THREADS, 0, st>>>(SA, CTX, N, T, intv, bwt32, tot, n1, N / FM_ROW + 1, reinterpret_cast<u32*>(occ2),
                                                        n2, b32, bocc_raw, pri);
  } else {
    const u64 rows = (N + FM_ROW - 1) / FM_ROW;
    k_fm_scan<<<(unsigned)rows, FM_ROW, 0, st>>>(SA, CTX, N, T, intv, bwt32, tot, n1, occ2, n2, b32, bocc_raw, pri);
  }
  CK_LAUNCH(); g_launches++;
}
void fm_occ1(const u32* tot, u64 stride, u64 n1, u32* occ1, cudaStream_t st) {
  k_fm_occ1<<<(unsigned)((n1 + 255) / 256), 256, 0, st>>>(tot, stride, n1, occ1);
  CK_LAUNCH(); g_launches++;
}
void fm_sample(const u32* SA, u64 N, u32 intv, const u32* b32, const u32* bocc, u32* sa_out, cudaStream_t st) {
  static const bool one = getenv("KISS_B200_FM_ONE") != nullptr;
  if ((((uintptr_t)SA) & 15) == 0 && !one)
    k_fm_sample4<<<(unsigned)((N + 1023) / 1024), 256, 0, st>>>(SA, N, intv, b32, bocc, sa_out);
  else
    k_fm_sample<<<(unsigned)((N + 255) / 256), 256, 0, st>>>(SA, N, intv, b32, bocc, sa_out);
  CK_LAUNCH(); g_launches++;
}

}  // namespace kiss
