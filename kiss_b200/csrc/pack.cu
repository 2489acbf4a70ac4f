// pack.cu — K1: text ingest into the 2-bit packed array, plus the ASCII coder and the synthetic-input
// generator.
//
// Replaces: PackedDNAString ctor (structs.hpp:84-116: 4 codes -> 1 byte, reversed layout) and, for the ASCII
// entry, Codec::to_int + `% 4` (include/biovoltron/utility/istring.hpp:27-46, include/command/suffix_sort.hpp:33).
// Layout: see common.cuh.  Streaming kernel: n bytes read, n/4 bytes written per launch.
#include "common.cuh"
#include "kernels.h"

namespace kiss {

u64 packed_words(u64 n) { return (n + 31) / 32 + KISS_PAD_WORDS; }

// 4 codes (one per byte, first base in the low byte) -> 8 bits, first base most significant
__device__ __forceinline__ u32 pack4(u32 x) {
  return ((x & 3u) << 6) | ((x >> 4) & 0x30u) | ((x >> 14) & 0x0Cu) | ((x >> 24) & 3u);
}

// one thread = 16 bases = one 32-bit half of a packed word
__global__ void __launch_bounds__(256) k_pack_codes(const u8* __restrict__ S, u64 n, u32* __restrict__ T32, u64 n16) {
  u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n16) return;
  u64 b = j * 16;
  u32 x, y, z, w;
  if (b + 16 <= n) {
    uint4 q = __ldg(reinterpret_cast<const uint4*>(S + b));
    x = q.x; y = q.y; z = q.z; w = q.w;
  } else {
    u32 v[4] = {0, 0, 0, 0};
    for (u32 t = 0; t < 16; t++)
      if (b + t < n) v[t >> 2] |= (u32)(S[b + t] & 3u) << (8 * (t & 3));
    x = v[0]; y = v[1]; z = v[2]; w = v[3];
  }
  u32 out = (pack4(x) << 24) | (pack4(y) << 16) | (pack4(z) << 8) | pack4(w);
  // the u64 word holds bases 0..15 in its high half: on a little-endian device that is the odd u32
  T32[j ^ 1] = out;
}

void pack_codes(const u8* S, u64 n, u64* T, cudaStream_t st) {
  u64 nw = packed_words(n);
  // zero the tail (the last text word may be half written, the pad words never are)
  u64 tail_from = n / 32;
  CK(cudaMemsetAsync(T + tail_from, 0, (nw - tail_from) * sizeof(u64), st));
  u64 n16 = (n + 15) / 16;
  if (n16 == 0) return;
  ProfScope ps(PK_PACK, st);
  k_pack_codes<<<(unsigned)((n16 + 255) / 256), 256, 0, st>>>(S, n, reinterpret_cast<u32*>(T), n16);
  CK_LAUNCH();
  g_launches++;
}

// ASCII -> code: aA 0, cC 1, gG 2, tT 3, anything else 4 % 4 = 0 (istring.hpp:28-36, suffix_sort.hpp:33)
__device__ __forceinline__ u32 code_of(u32 c) {
  u32 up = c & 0xDFu;
  u32 v = ((c >> 1) ^ (c >> 2)) & 3u;
  bool ok = (up == 'A') | (up == 'C') | (up == 'G') | (up == 'T');
  return ok ? v : 0u;
}

__global__ void __launch_bounds__(256) k_ascii_to_codes(const u8* __restrict__ a, u64 n, u8* __restrict__ S) {
  u64 j = ((u64)blockIdx.x * blockDim.x + threadIdx.x) * 16;
  if (j >= n) return;
  if (j + 16 <= n) {
    uint4 q = __ldg(reinterpret_cast<const uint4*>(a + j));
    u32 in[4] = {q.x, q.y, q.z, q.w}, out[4];
#pragma unroll
    for (int t = 0; t < 4; t++) {
      u32 x = in[t];
      out[t] = code_of(x & 255u) | (code_of((x >> 8) & 255u) << 8) | (code_of((x >> 16) & 255u) << 16) |
               (code_of(x >> 24) << 24);
    }
    *reinterpret_cast<uint4*>(S + j) = make_uint4(out[0], out[1], out[2], out[3]);
  } else {
    for (u64 t = j; t < n; t++) S[t] = (u8)code_of(a[t]);
  }
}

void ascii_to_codes(const u8* ascii, u64 n, u8* S, cudaStream_t st) {
  if (n == 0) return;
  u64 nt = (n + 15) / 16;
  k_ascii_to_codes<<<(unsigned)((nt + 255) / 256), 256, 0, st>>>(ascii, n, S);
  CK_LAUNCH();
  g_launches++;
}

// base i = (splitmix64(seed' ^ (i >> 5)) >> 2(i & 31)) & 3, seed' = splitmix64(seed)
__global__ void __launch_bounds__(256) k_synth_uniform(u8* __restrict__ S, u64 n, u64 s) {
  u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x;  // one thread = 16 bases
  u64 b = j * 16;
  if (b >= n) return;
  u64 x = splitmix64(s ^ (b >> 5));
  u32 bits = (u32)(x >> (2 * (b & 31)));  // 16 bases, first in the low bits
  u32 out[4];
#pragma unroll
  for (int t = 0; t < 4; t++) {
    u32 q = (bits >> (8 * t)) & 255u;
    out[t] = (q & 3u) | (((q >> 2) & 3u) << 8) | (((q >> 4) & 3u) << 16) | (((q >> 6) & 3u) << 24);
  }
  if (b + 16 <= n) {
    *reinterpret_cast<uint4*>(S + b) = make_uint4(out[0], out[1], out[2], out[3]);
  } else {
    for (u32 t = 0; b + t < n; t++) S[b + t] = (u8)((out[t >> 2] >> (8 * (t & 3))) & 3u);
  }
}

void synth_uniform(u8* S, u64 n, u64 seed, cudaStream_t st) {
  if (n == 0) return;
  u64 nt = (n + 15) / 16;
  k_synth_uniform<<<(unsigned)((nt + 255) / 256), 256, 0, st>>>(S, n, splitmix64(seed));
  CK_LAUNCH();
  g_launches++;
}

// Position-mixed checksum of an array of suffix indices: sum_i splitmix64((i << 32) | a[i]) mod 2^64 — sensitive to
// the order of the entries, yet a plain sum, so it is one streaming pass (tests/golden/make_full_size_golden.py computes
// the same number from the reference's array).  *out must be zeroed by the caller.
__global__ void __launch_bounds__(256) k_mixed_sum(const u32* __restrict__ a, u64 count, u64* __restrict__ out) {
  u64 acc = 0;
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (u64)gridDim.x * blockDim.x)
    acc += splitmix64((i << 32) | (u64)a[i]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31u) == 0) atomicAdd((unsigned long long*)out, (unsigned long long)acc);
}
void mixed_sum(const u32* a, u64 count, u64* out_dev, cudaStream_t st) {
  CK(cudaMemsetAsync(out_dev, 0, sizeof(u64), st));
  if (!count) return;
  k_mixed_sum<<<148 * 8, 256, 0, st>>>(a, count, out_dev);
  CK_LAUNCH();
  g_launches++;
}

}  // namespace kiss
