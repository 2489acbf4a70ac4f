// verify.cu — the property the reference's own tests assert (tests/kiss.cpp:25-33, 61-65), checked on the device over the
// WHOLE array: SA[0] = n, SA is a permutation of 0..n, and substr(SA[i-1], k) <= substr(SA[i], k) for every i (the shorter
// string first when one is a prefix of the other).  Size-independent, so bench.py runs it at the full 3.1 Gbp next to the
// digest comparison with the reference's array; no counterpart in the reference's library (its check is the test's loop).
#include "common.cuh"
#include "kernels.h"

namespace kiss {

// out[0] = ordered-pair violations, out[1] = values seen twice, out[2] = values out of range or SA[0] != n,
// out[3] = pairs equal over `cap_words` words whose comparison was cut there (only with huge k on repetitive text)
__global__ void __launch_bounds__(256) k_verify_sa(const u64* __restrict__ T, u64 n, const u32* __restrict__ SA, u64 N, u64 k,
                                                  u32 cap_words, u32* __restrict__ bitmap, u64* __restrict__ out) {
  const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const u64 b = SA[i];
  if (b > n || (i == 0 && b != n)) {
    atomicAdd(&out[2], 1ull);
    return;
  }
  const u32 bit = 1u << (b & 31);
  if (atomicOr(&bitmap[b >> 5], bit) & bit) atomicAdd(&out[1], 1ull);
  if (i == 0) return;
  const u64 a = SA[i - 1];
  if (a > n) return;  // (counted by its own thread)
  const u64 la = n - a, lb = n - b;
  u64 L = la < lb ? la : lb;
  if (k < L) L = k;
  u32 it = 0;
  for (u64 q = 0; q < L; q += 32, it++) {
    if (it >= cap_words) { atomicAdd(&out[3], 1ull); return; }
    u64 wa = word_at(T, a + q), wb = word_at(T, b + q);
    const u64 rem = L - q;
    if (rem < 32) { const u64 mk = ~0ull << (64 - 2 * rem); wa &= mk; wb &= mk; }
    if (wa != wb) {
      if (wa > wb) atomicAdd(&out[0], 1ull);
      return;
    }
  }
  if (L < k && la > lb) atomicAdd(&out[0], 1ull);  // equal as far as the shorter goes: the shorter must come first
}

// the same over a byte text (general alphabet): symbols compared one by one, `cap` symbols at most
__global__ void __launch_bounds__(256) k_verify_sa_bytes(const u8* __restrict__ S, u64 n, const u32* __restrict__ SA, u64 N, u64 k,
                                                        u32 cap, u32* __restrict__ bitmap, u64* __restrict__ out) {
  const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const u64 b = SA[i];
  if (b > n || (i == 0 && b != n)) {
    atomicAdd(&out[2], 1ull);
    return;
  }
  const u32 bit = 1u << (b & 31);
  if (atomicOr(&bitmap[b >> 5], bit) & bit) atomicAdd(&out[1], 1ull);
  if (i == 0) return;
  const u64 a = SA[i - 1];
  if (a > n) return;
  const u64 la = n - a, lb = n - b;
  u64 L = la < lb ? la : lb;
  if (k < L) L = k;
  for (u64 q = 0; q < L; q++) {
    if (q >= cap) { atomicAdd(&out[3], 1ull); return; }
    const u8 ca = S[a + q], cb = S[b + q];
    if (ca != cb) {
      if (ca > cb) atomicAdd(&out[0], 1ull);
      return;
    }
  }
  if (L < k && la > lb) atomicAdd(&out[0], 1ull);
}

void verify_sa_bytes(const u8* S, u64 n, const u32* SA, u64 N, u64 k, u32 cap, u32* bitmap, u64* out, cudaStream_t st) {
  CK(cudaMemsetAsync(bitmap, 0, ((N + 31) / 32) * sizeof(u32), st));
  CK(cudaMemsetAsync(out, 0, 4 * sizeof(u64), st));
  if (!N) return;
  k_verify_sa_bytes<<<(unsigned)((N + 255) / 256), 256, 0, st>>>(S, n, SA, N, k, cap, bitmap, out);
  CK_LAUNCH(); g_launches++;
}

void verify_sa(const u64* T, u64 n, const u32* SA, u64 N, u64 k, u32 cap_words, u32* bitmap, u64* out, cudaStream_t st) {
  CK(cudaMemsetAsync(bitmap, 0, ((N + 31) / 32) * sizeof(u32), st));
  CK(cudaMemsetAsync(out, 0, 4 * sizeof(u64), st));
  if (!N) return;
  k_verify_sa<<<(unsigned)((N + 255) / 256), 256, 0, st>>>(T, n, SA, N, k, cap_words, bitmap, out);
  CK_LAUNCH(); g_launches++;
}

}  // namespace kiss
