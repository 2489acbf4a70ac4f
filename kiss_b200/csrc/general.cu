// general.cu — the general-alphabet entry: suffix array of a text over any byte alphabet.
//
// Replaces KISS1Sorter<uint32_t>::get_suffix_array (kiss1_sorter.hpp:28-44) -> kiss::kiss1_suffix_array
// (kiss1_core.hpp:270-311: get_lms, lms_suffix_direct_sort :153-226, put_lms_suffix, induced_sort over 256 buckets) and
// the same call of KISS2Sorter.
//
// The reference runs its pipeline a second time over bytes.  Here the text is REDUCED to the 2-bit alphabet the device
// path is built for: the sigma distinct byte values of the text are replaced by their ranks, every rank is written as
// b = max(1, ceil(log4 sigma)) two-bit digits, most significant first, and the resulting DNA text S' of b*n bases goes
// through the unchanged DNA pipeline.  Fixed-width, order-preserving codes make the comparison of the suffixes b*i and
// b*j of S' the comparison of the suffixes i and j of S, the end of the text included (the shorter one first), so the
// suffix array of S is the sub-sequence of the entries of SA(S') that are multiples of b, divided by b.  For sigma <= 4
// (the reference's own tests use 'A'..'D', tests/kiss.cpp:53-87) b = 1 and nothing is filtered.
//
// Order: the DNA path is run with k' = b * k, which orders every suffix of S' by at least its first b*k bases, i.e.
// every suffix of S by at least its first k symbols — the property the reference's tests assert (tests/kiss.cpp:61-65).
// Where the reference's result is unique (no two suffixes share their first k symbols, or k >= n - 1: the true suffix
// array) the result is identical to it; among suffixes that DO share k symbols the reference's order is whatever its
// unstable std::sort leaves for n < 10^8 (kiss1_core.hpp:214-219) and "32-symbol blocks, then position" above, and the
// order here is the DNA path's (deeper than k, then induced) — a valid k-ordered suffix array, not the same permutation.
#include <algorithm>
#include <string>

#include "../../include/kiss_b200.h"
#include "common.cuh"
#include "kernels.h"

using namespace kiss;

namespace {

#define GEN_THREADS 256
#define GEN_ITEMS 16
#define GEN_TILE (GEN_THREADS * GEN_ITEMS)

// which byte values occur
__global__ void __launch_bounds__(GEN_THREADS) k_gen_presence(const u8* __restrict__ S, u64 n, u32* __restrict__ present) {
  __shared__ u32 s[256];
  s[threadIdx.x] = 0;
  __syncthreads();
  const u64 stride = (u64)gridDim.x * GEN_THREADS * 16;
  for (u64 i = ((u64)blockIdx.x * GEN_THREADS + threadIdx.x) * 16; i < n; i += stride) {
    if (i + 16 <= n && (((uintptr_t)S) & 15) == 0) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(S + i));
      const u32 w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int q = 0; q < 4; q++)
#pragma unroll
        for (int e = 0; e < 4; e++) s[(w[q] >> (8 * e)) & 0xFFu] = 1u;  // (same value from every writer)
    } else {
      for (u64 j = i; j < n && j < i + 16; j++) s[S[j]] = 1u;
    }
  }
  __syncthreads();
  if (s[threadIdx.x]) present[threadIdx.x] = 1u;
}

// rank codes as b two-bit digits per symbol, most significant first: one code byte per base, the input format of the
// DNA entry (its k_pack_codes folds them into the packed text)
template <int B>
__global__ void __launch_bounds__(GEN_THREADS) k_gen_expand(const u8* __restrict__ S, u64 n, const u8* __restrict__ rank,
                                                           u8* __restrict__ codes) {
  __shared__ u8 r[256];
  r[threadIdx.x] = rank[threadIdx.x];
  __syncthreads();
  const u64 stride = (u64)gridDim.x * GEN_THREADS;
  for (u64 i = (u64)blockIdx.x * GEN_THREADS + threadIdx.x; i < n; i += stride) {
    const u32 c = r[S[i]];
    if (B == 1) {
      codes[i] = (u8)c;
    } else if (B == 2) {
      reinterpret_cast<unsigned short*>(codes)[i] = (unsigned short)(((c >> 2) & 3u) | ((c & 3u) << 8));
    } else if (B == 3) {
      codes[3 * i] = (u8)((c >> 4) & 3u);
      codes[3 * i + 1] = (u8)((c >> 2) & 3u);
      codes[3 * i + 2] = (u8)(c & 3u);
    } else {
      reinterpret_cast<u32*>(codes)[i] = ((c >> 6) & 3u) | (((c >> 4) & 3u) << 8) | (((c >> 2) & 3u) << 16) | ((c & 3u) << 24);
    }
  }
}

// entries of SA(S') at symbol boundaries, per tile of 4096
__global__ void __launch_bounds__(GEN_THREADS) k_gen_count(const u32* __restrict__ SAp, u64 N, u32 b, u32* __restrict__ cnt) {
  __shared__ u32 ws[33];
  const u64 e0 = (u64)blockIdx.x * GEN_TILE + (u64)threadIdx.x * GEN_ITEMS;
  u32 c = 0;
#pragma unroll
  for (int i = 0; i < GEN_ITEMS; i++)
    if (e0 + i < N && SAp[e0 + i] % b == 0) c++;
  u32 total;
  block_excl_scan(c, ws, total);
  if (threadIdx.x == 0) cnt[blockIdx.x] = total;
}

__global__ void __launch_bounds__(GEN_THREADS) k_gen_emit(const u32* __restrict__ SAp, u64 N, u32 b, const u32* __restrict__ off,
                                                         u32* __restrict__ SA) {
  __shared__ u32 ws[33];
  const u64 e0 = (u64)blockIdx.x * GEN_TILE + (u64)threadIdx.x * GEN_ITEMS;
  u32 v[GEN_ITEMS];
  u32 c = 0;
#pragma unroll
  for (int i = 0; i < GEN_ITEMS; i++) {
    v[i] = e0 + i < N ? SAp[e0 + i] : 1u;
    if (e0 + i < N && v[i] % b == 0) c++;
  }
  u32 total;
  u64 o = (u64)off[blockIdx.x] + block_excl_scan(c, ws, total);
#pragma unroll
  for (int i = 0; i < GEN_ITEMS; i++)
    if (e0 + i < N && v[i] % b == 0) SA[o++] = v[i] / b;
}

struct DevBuf {
  void* p = nullptr;
  cudaStream_t st;
  explicit DevBuf(cudaStream_t s) : st(s) {}
  void alloc(u64 bytes) { CK(cudaMallocAsync(&p, bytes ? bytes : 1, st)); }
  ~DevBuf() { if (p) cudaFreeAsync(p, st); }
  template <typename T> T* as() { return static_cast<T*>(p); }
};

// the distinct byte values of the text: rank[v] = number of smaller values that occur; sigma of them; b digits each
void alphabet_of(const u8* S_dev, u64 n, cudaStream_t st, u8 rank[256], u32* sigma, u32* b) {
  DevBuf present(st);
  present.alloc(256 * sizeof(u32));
  CK(cudaMemsetAsync(present.p, 0, 256 * sizeof(u32), st));
  const unsigned grid = (unsigned)std::min<u64>((n + GEN_TILE - 1) / GEN_TILE, 148ull * 8);
  k_gen_presence<<<grid, GEN_THREADS, 0, st>>>(S_dev, n, present.as<u32>());
  CK_LAUNCH(); g_launches++;
  u32 pres[256];
  CK(cudaMemcpyAsync(pres, present.p, sizeof(pres), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  u32 s = 0;
  for (int v = 0; v < 256; v++) {
    rank[v] = (u8)s;
    if (pres[v]) s++;
  }
  u32 d = 1;
  while (d < 4 && (1u << (2 * d)) < s) d++;
  *sigma = s;
  *b = d;
}

int general_dev(kiss_b200_ctx* c, const u8* S_dev, u64 n, u32 k, int algo, u32* SA_dev, kiss_b200_timings* t) {
  const int device = api_ctx_device(c);
  CK(cudaSetDevice(device));
  cudaStream_t st = static_cast<cudaStream_t>(kiss_b200_stream(c));
  g_prof = nullptr;
  DevBuf rank_d(st);
  rank_d.alloc(256);
  u8 rank[256];
  u32 sigma = 0, b = 1;
  alphabet_of(S_dev, n, st, rank, &sigma, &b);
  const u64 np = n * b;  // bases of the reduced text
  if (np + 1 >= 0xFFFFFFFFull)
    throw CudaError(KISS_B200_ERR_TOO_LARGE, "general alphabet: " + std::to_string(n) + " symbols of " + std::to_string(sigma) +
                    " kinds need " + std::to_string(b) + " bases each, beyond 32-bit suffix indices");
  CK(cudaMemcpyAsync(rank_d.p, rank, 256, cudaMemcpyHostToDevice, st));
  DevBuf codes(st);
  codes.alloc(np + 16);
  const unsigned g2 = (unsigned)std::min<u64>((n + GEN_THREADS - 1) / GEN_THREADS, 148ull * 16);
  switch (b) {
    case 1: k_gen_expand<1><<<g2, GEN_THREADS, 0, st>>>(S_dev, n, rank_d.as<u8>(), codes.as<u8>()); break;
    case 2: k_gen_expand<2><<<g2, GEN_THREADS, 0, st>>>(S_dev, n, rank_d.as<u8>(), codes.as<u8>()); break;
    case 3: k_gen_expand<3><<<g2, GEN_THREADS, 0, st>>>(S_dev, n, rank_d.as<u8>(), codes.as<u8>()); break;
    default: k_gen_expand<4><<<g2, GEN_THREADS, 0, st>>>(S_dev, n, rank_d.as<u8>(), codes.as<u8>()); break;
  }
  CK_LAUNCH(); g_launches++;
  CK(cudaStreamSynchronize(st));  // (rank[] is on this frame)
  // k symbols = b * k bases; k >= n - 1 asks for the true suffix array
  const u64 kp64 = (u64)k * b;
  const u32 kp = ((u64)k + 1 >= n || kp64 >= 0xFFFFFFFFull) ? 0xFFFFFFFFu : (u32)kp64;
  u64 mine = 2;  // presence, expand
  int rc;
  if (b == 1) {
    rc = kiss_b200_suffix_sort_dna_u32_dev(c, codes.as<u8>(), np, kp, algo, SA_dev, t);
    if (rc != KISS_B200_OK) return rc;
  } else {
    DevBuf sap(st), cnt(st), part(st);
    const u64 N = np + 1;
    sap.alloc(N * sizeof(u32));
    rc = kiss_b200_suffix_sort_dna_u32_dev(c, codes.as<u8>(), np, kp, algo, sap.as<u32>(), t);
    if (rc != KISS_B200_OK) return rc;
    g_prof = nullptr;  // (the inner call pointed it at the context's profiler: the scan below is not part of a profiled call)
    const u64 tiles = (N + GEN_TILE - 1) / GEN_TILE;
    cnt.alloc(tiles * sizeof(u32));
    part.alloc(scan_partials_len(tiles) * sizeof(u32));
    k_gen_count<<<(unsigned)tiles, GEN_THREADS, 0, st>>>(sap.as<u32>(), N, b, cnt.as<u32>());
    CK_LAUNCH(); g_launches++;
    scan_excl_u32(cnt.as<u32>(), tiles, part.as<u32>(), nullptr, st);
    k_gen_emit<<<(unsigned)tiles, GEN_THREADS, 0, st>>>(sap.as<u32>(), N, b, cnt.as<u32>(), SA_dev);
    CK_LAUNCH(); g_launches++;
    CK(cudaStreamSynchronize(st));
    mine += 5;  // count, three scan kernels, emit
  }
  if (t) {
    t->n = n;  // (m, the wave counts and the stage times are those of the reduced text)
    t->kernel_launches += mine;
  }
  return KISS_B200_OK;
}

}  // namespace

extern "C" int kiss_b200_suffix_sort_u32_dev(kiss_b200_ctx* c, const uint8_t* S_dev, uint64_t n, uint32_t k, int algo,
                                             uint32_t* SA_dev, kiss_b200_timings* t) {
  if (!c || !SA_dev || (n && !S_dev)) {
    api_set_last_error("bad argument");
    return KISS_B200_ERR_ARG;
  }
  try {
    if (n == 0) return kiss_b200_suffix_sort_dna_u32_dev(c, S_dev, 0, k, algo, SA_dev, t);  // {0} (kiss1_core.hpp:278-279)
    return general_dev(c, S_dev, n, k, algo, SA_dev, t);
  } catch (const CudaError& e) {
    api_set_last_error(e.what());
    cudaGetLastError();
    return e.code;
  }
}

extern "C" int kiss_b200_suffix_sort_u32(kiss_b200_ctx* c, const uint8_t* S_host, uint64_t n, uint32_t k, int algo,
                                         uint32_t* SA_host, kiss_b200_timings* t) {
  if (!c || !SA_host || (n && !S_host)) {
    api_set_last_error("bad argument");
    return KISS_B200_ERR_ARG;
  }
  if (n == 0) {
    SA_host[0] = 0;
    if (t) *t = kiss_b200_timings{};
    return KISS_B200_OK;
  }
  try {
    CK(cudaSetDevice(api_ctx_device(c)));
    cudaStream_t st = static_cast<cudaStream_t>(kiss_b200_stream(c));
    DevBuf S(st), SA(st);
    S.alloc(n + 16);
    SA.alloc((n + 1) * sizeof(u32));
    cudaEvent_t e0, e1, e2, e3;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1)); CK(cudaEventCreate(&e2)); CK(cudaEventCreate(&e3));
    struct Ev { cudaEvent_t a, b, c, d; ~Ev() { cudaEventDestroy(a); cudaEventDestroy(b); cudaEventDestroy(c); cudaEventDestroy(d); } } ev{e0, e1, e2, e3};
    CK(cudaEventRecord(e0, st));
    CK(cudaMemcpyAsync(S.p, S_host, n, cudaMemcpyHostToDevice, st));
    CK(cudaEventRecord(e1, st));
    const int rc = general_dev(c, S.as<u8>(), n, k, algo, SA.as<u32>(), t);
    if (rc != KISS_B200_OK) return rc;
    CK(cudaEventRecord(e2, st));
    CK(cudaMemcpyAsync(SA_host, SA.p, (n + 1) * sizeof(u32), cudaMemcpyDeviceToHost, st));
    CK(cudaEventRecord(e3, st));
    CK(cudaStreamSynchronize(st));
    if (t) {
      float a = 0, b = 0;
      CK(cudaEventElapsedTime(&a, e0, e1));
      CK(cudaEventElapsedTime(&b, e2, e3));
      t->h2d = a * 1e-3;
      t->d2h = b * 1e-3;
    }
    return KISS_B200_OK;
  } catch (const CudaError& e) {
    api_set_last_error(e.what());
    cudaGetLastError();
    return e.code;
  }
}

extern "C" int kiss_b200_general_digits(kiss_b200_ctx* c, const uint8_t* S_dev, uint64_t n, uint32_t* sigma, uint32_t* digits) {
  if (!c || !sigma || !digits || (n && !S_dev)) {
    api_set_last_error("bad argument");
    return KISS_B200_ERR_ARG;
  }
  *sigma = 0;
  *digits = 1;
  if (!n) return KISS_B200_OK;
  try {
    CK(cudaSetDevice(api_ctx_device(c)));
    cudaStream_t st = static_cast<cudaStream_t>(kiss_b200_stream(c));
    u8 rank[256];
    u32 s = 0, b = 1;
    alphabet_of(S_dev, n, st, rank, &s, &b);
    *sigma = s;
    *digits = b;
    return KISS_B200_OK;
  } catch (const CudaError& e) {
    api_set_last_error(e.what());
    cudaGetLastError();
    return e.code;
  }
}
