// common.cuh — shared device helpers of the B200 suffix_sort path.
//
// Packed text layout in HBM (differs from the reference's reversed DibitVector, structs.hpp:84-116;
// SURVEY.md A.1): 32 bases per 64-bit word, base i in word i/32 at bits [62-2(i%32), 63-2(i%32)],
// i.e. the first base of a word is its most significant field, so an integer compare of words is a
// lexicographic compare of 32-mers.  Words past the text are zero ('A'), like the reference's padding
// (structs.hpp:94-96).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

typedef unsigned long long u64;
typedef unsigned int u32;
typedef unsigned char u8;

#define KISS_EMPTY 0xFFFFFFFFu
#define KISS_PAD_WORDS 4  // zero words kept after the last text word

#define FIELD_LO 0x5555555555555555ull

__device__ __forceinline__ u32 lane_id() { return threadIdx.x & 31; }

// base at text position i (i < n + 32*KISS_PAD_WORDS)
__device__ __forceinline__ u32 base_at(const u64* __restrict__ T, u64 i) {
  return (u32)(__ldg(T + (i >> 5)) >> (62 - 2 * (i & 31))) & 3u;
}

// the 32 bases starting at text position p as one word (first base most significant);
// bases past the end of the text read as 0.  Reference: load_prefix_length_125 / _less_than_16
// (structs.hpp:122-184) read the same big-endian k-mers from the reversed layout.
__device__ __forceinline__ u64 word_at(const u64* __restrict__ T, u64 p) {
  u64 i = p >> 5;
  u32 s = (u32)(p & 31) * 2;
  u64 a = __ldg(T + i);
  if (s == 0) return a;
  u64 b = __ldg(T + i + 1);
  return (a << s) | (b >> (64 - s));
}

// ---- field-wise (2-bit) comparisons; results have one bit per base at the low bit of its field
__device__ __forceinline__ void field_cmp(u64 x, u64 y, u64& lt, u64& eq) {
  u64 ah = (x >> 1) & FIELD_LO, al = x & FIELD_LO, bh = (y >> 1) & FIELD_LO, bl = y & FIELD_LO;
  u64 heq = ~(ah ^ bh) & FIELD_LO;
  lt = (~ah & bh & FIELD_LO) | (heq & ~al & bl);
  eq = heq & ~(al ^ bl);
}

// mask with the low field bit set for bases [0, cnt) of a word (cnt in 0..32), base 0 = top field
__device__ __forceinline__ u64 first_fields(u32 cnt) {
  if (cnt >= 32) return FIELD_LO;
  if (cnt == 0) return 0;
  return FIELD_LO & ~((1ull << (64 - 2 * cnt)) - 1ull);
}

// ---- warp / block scans -------------------------------------------------------------------------
__device__ __forceinline__ u32 warp_incl_scan(u32 v) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    u32 t = __shfl_up_sync(0xffffffffu, v, o);
    if ((int)lane_id() >= o) v += t;
  }
  return v;
}

// exclusive scan over the blockDim.x (<= 1024, multiple of 32) values `v`; returns the exclusive
// prefix of this thread and the block total in `total`.  `ws` is 33 u32 of shared scratch.
__device__ __forceinline__ u32 block_excl_scan(u32 v, u32* ws, u32& total) {
  u32 inc = warp_incl_scan(v);
  u32 w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  __syncthreads();
  if (lane_id() == 31) ws[w] = inc;
  __syncthreads();
  if (w == 0) {
    u32 x = lane_id() < nw ? ws[lane_id()] : 0;
    u32 xi = warp_incl_scan(x);
    ws[lane_id()] = xi - x;
    if (lane_id() == 31) ws[32] = xi;
  }
  __syncthreads();
  total = ws[32];
  return ws[w] + inc - v;
}

// splitmix64 — the counter hash of the synthetic inputs (tests/textgen.py)
__host__ __device__ __forceinline__ u64 splitmix64(u64 z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
