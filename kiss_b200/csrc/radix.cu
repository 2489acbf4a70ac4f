// radix.cu — device-wide exclusive scan and the stable LSD radix sort of (key, value) pairs used by the
// k-ordered LMS sort (replaces the reference's 2^20-bucket counting sort + per-bucket std::sort,
// kiss1_core.hpp:41-83, 137-142).  Hand-written for sm_100a; no CUB.
//
// One radix pass = digit histogram per 4096-key tile (k_radix_hist), one exclusive scan over the
// digit-major [256][tiles] counter matrix, and a scatter (k_radix_scatter) that ranks the tile's keys
// stably with warp match/ballot, stages the tile in shared memory in digit order and writes each digit's
// run with coalesced stores.  Algorithmic bytes per pass and key: sizeof(K) (histogram read)
// + 2 * (sizeof(K) + 4) (scatter read + write).
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"

namespace kiss {

thread_local u64 g_launches = 0;
thread_local Profiler* g_prof = nullptr;

const char* prof_name(int id) {
  static const char* names[PK_COUNT] = {"k_pack_codes", "k_lms_summary+carry+mask", "k_lms_emit", "k_scan_*", "k_radix_hist",
                                        "k_radix_scatter", "k_groups_tiny", "k_groups_medium<=32", "k_groups_medium<=64",
                                        "k_groups_medium<=128", "k_groups_medium<=256", "k_groups_medium<=512",
                                        "k_groups_medium<=2048", "k_groups_classify+big_*",
                                        "k_short_*", "k_place_lms", "k_wave_fused(lms)", "k_wave_fused", "k_wave_small/tail"};
  return (id >= 0 && id < PK_COUNT) ? names[id] : "?";
}
cudaEvent_t Profiler::get() {
  if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
  cudaEvent_t e;
  cudaEventCreate(&e);
  return e;
}
void Profiler::collect() {
  for (auto& r : recs) {
    float t = 0;
    if (cudaEventElapsedTime(&t, r.a, r.b) == cudaSuccess) { ms[r.id] += t; launches[r.id]++; }
    pool.push_back(r.a); pool.push_back(r.b);
  }
  recs.clear();
}
void Profiler::reset() {
  for (int i = 0; i < PK_COUNT; i++) { ms[i] = 0; launches[i] = 0; bytes[i] = 0; }
}
Profiler::~Profiler() {
  collect();
  for (auto e : pool) cudaEventDestroy(e);
}

void cuda_check(cudaError_t e, const char* what, const char* file, int line) {
  if (e == cudaSuccess) return;
  std::string m = std::string(cudaGetErrorName(e)) + ": " + cudaGetErrorString(e) + " at " + file + ":" +
                  std::to_string(line) + " (" + what + ")";
  int code = (e == cudaErrorMemoryAllocation) ? -4 : -3;
  throw CudaError(code, m);
}

// =================================================================================== scan
#define SCAN_THREADS 256
#define SCAN_ITEMS 16
#define SCAN_TILE 4096

__global__ void __launch_bounds__(SCAN_THREADS) k_scan_reduce(const u32* __restrict__ in, u64 N, u32* __restrict__ partials) {
  __shared__ u32 ws[33];
  u64 base = (u64)blockIdx.x * SCAN_TILE;
  u32 s = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; i++) {
    u64 idx = base + (u64)i * SCAN_THREADS + threadIdx.x;
    if (idx < N) s += in[idx];
  }
  u32 tot;
  block_excl_scan(s, ws, tot);
  if (threadIdx.x == 0) partials[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(1024) k_scan_partials(u32* __restrict__ partials, u32 P, u32* __restrict__ total_out) {
  __shared__ u32 ws[33];
  u32 carry = 0;
  for (u32 base = 0; base < P; base += 1024) {
    u32 idx = base + threadIdx.x;
    u32 v = idx < P ? partials[idx] : 0;
    u32 tot;
    u32 ex = block_excl_scan(v, ws, tot);
    if (idx < P) partials[idx] = carry + ex;
    carry += tot;
  }
  if (threadIdx.x == 0 && total_out) *total_out = carry;
}

// (`in` == `data` for the in-place scan: no __restrict__ on the two)
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_down(const u32* in, u32* data, u64 N, const u32* __restrict__ partials) {
  __shared__ u32 ws[33];
  u64 base = (u64)blockIdx.x * SCAN_TILE + (u64)threadIdx.x * SCAN_ITEMS;
  u32 v[SCAN_ITEMS];
  bool full = base + SCAN_ITEMS <= N;
  if (full) {
    const uint4* p = reinterpret_cast<const uint4*>(in + base);
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS / 4; i++) {
      uint4 q = p[i];
      v[4 * i] = q.x; v[4 * i + 1] = q.y; v[4 * i + 2] = q.z; v[4 * i + 3] = q.w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) v[i] = (base + i < N) ? in[base + i] : 0;
  }
  u32 s = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; i++) { u32 t = v[i]; v[i] = s; s += t; }
  u32 tot;
  u32 ex = block_excl_scan(s, ws, tot) + partials[blockIdx.x];
  if (full) {
    uint4* p = reinterpret_cast<uint4*>(data + base);
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS / 4; i++)
      p[i] = make_uint4(v[4 * i] + ex, v[4 * i + 1] + ex, v[4 * i + 2] + ex, v[4 * i + 3] + ex);
  } else {
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++)
      if (base + i < N) data[base + i] = v[i] + ex;
  }
}

// out = exclusive prefix sums of in (in == out: in place); both 16-byte aligned
void scan_excl_u32_to(const u32* in, u32* out, u64 N, u32* partials, u32* total_dev, cudaStream_t st) {
  if (N == 0) {
    if (total_dev) CK(cudaMemsetAsync(total_dev, 0, sizeof(u32), st));
    return;
  }
  ProfScope ps(PK_SCAN, st);
  u32 P = (u32)((N + SCAN_TILE - 1) / SCAN_TILE);
  k_scan_reduce<<<P, SCAN_THREADS, 0, st>>>(in, N, partials);
  CK_LAUNCH();
  k_scan_partials<<<1, 1024, 0, st>>>(partials, P, total_dev);
  CK_LAUNCH();
  k_scan_down<<<P, SCAN_THREADS, 0, st>>>(in, out, N, partials);
  CK_LAUNCH();
  g_launches += 3;
}
void scan_excl_u32(u32* data, u64 N, u32* partials, u32* total_dev, cudaStream_t st) {
  scan_excl_u32_to(data, data, N, partials, total_dev, st);
}

// =================================================================================== radix
#ifndef RADIX_THREADS
#define RADIX_THREADS 256  // 512 threads x 8 keys (2 blocks/SM, 32 warps) measured slower: scatter 7.86 vs 7.50 ms, hist 1.24 vs 1.00 ms
#endif
#define RADIX_TILE 4096
#define RADIX_ITEMS (RADIX_TILE / RADIX_THREADS)
#define RADIX_WARPS (RADIX_THREADS / 32)
#define RADIX_MINB32 (RADIX_THREADS == 256 ? 3 : 2)

template <typename K>
__global__ void __launch_bounds__(RADIX_THREADS) k_radix_hist(const K* __restrict__ kin, u32 N, int shift, u32 mask,
                                                             u32* __restrict__ H, u32 ntiles) {
  __shared__ u32 hist[RADIX_WARPS][256];  // one sub-histogram per warp keeps shared-memory atomics uncontended
  const u32 tid = threadIdx.x, tile = blockIdx.x, w = tid >> 5;
  for (u32 i = tid; i < RADIX_WARPS * 256; i += RADIX_THREADS) (&hist[0][0])[i] = 0;
  __syncthreads();
  const u32 tile_base = tile * RADIX_TILE;
  K keys[RADIX_ITEMS];
#pragma unroll
  for (int i = 0; i < RADIX_ITEMS; i++) {
    u32 e = tile_base + i * RADIX_THREADS + tid;
    keys[i] = e < N ? kin[e] : (K)0;
  }
#pragma unroll
  for (int i = 0; i < RADIX_ITEMS; i++) {
    u32 e = tile_base + i * RADIX_THREADS + tid;
    if (e < N) atomicAdd(&hist[w][(u32)(keys[i] >> shift) & mask], 1u);
  }
  __syncthreads();
  if (tid < 256) {
    u32 t = 0;
#pragma unroll
    for (int ww = 0; ww < RADIX_WARPS; ww++) t += hist[ww][tid];
    H[(u64)tid * ntiles + tile] = t;
  }
}

template <typename K>
__global__ void __launch_bounds__(RADIX_THREADS, RADIX_MINB32) k_radix_scatter(const K* __restrict__ kin, const u32* __restrict__ vin,
                                                                K* __restrict__ kout, u32* __restrict__ vout, u32 N,
                                                                int shift, u32 mask, const u32* __restrict__ H,
                                                                u32 ntiles) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  K* skeys = reinterpret_cast<K*>(smem_raw);
  u32* svals = reinterpret_cast<u32*>(skeys + RADIX_TILE);
  u32* wcnt = svals + RADIX_TILE;          // [RADIX_WARPS][256]
  u32* texcl = wcnt + RADIX_WARPS * 256;   // [256] exclusive digit offsets inside the tile
  u32* dbase = texcl + 256;                // [256] global base of the digit run minus texcl
  u32* ws = dbase + 256;                   // [33]

  const u32 tid = threadIdx.x, tile = blockIdx.x, w = tid >> 5, lane = tid & 31;
  const u32 tile_base = tile * RADIX_TILE;
  const u32 tile_n = min((u32)RADIX_TILE, N - tile_base);
  for (u32 i = tid; i < RADIX_WARPS * 256; i += RADIX_THREADS) wcnt[i] = 0;
  __syncthreads();

  K keys[RADIX_ITEMS];
  u32 rank[RADIX_ITEMS];
  const u32 wbase = tile_base + w * (32 * RADIX_ITEMS);
  const u32 lt_mask = (1u << lane) - 1u;
  // keys and this thread's digit base are requested up front; values are fetched when staged (holding them
  // in registers too costs a resident block per SM, which measured slower)
#pragma unroll
  for (int r = 0; r < RADIX_ITEMS; r++) {
    u32 e = wbase + r * 32 + lane;
    keys[r] = e < N ? kin[e] : (K)0;
  }
  const u32 hbase = tid < 256 ? H[(u64)tid * ntiles + tile] : 0u;
#pragma unroll
  for (int r = 0; r < RADIX_ITEMS; r++) {
    u32 e = wbase + r * 32 + lane;
    bool valid = e < N;
    K key = keys[r];
    u32 d = (u32)(key >> shift) & mask;
    // lanes holding the same digit: one ballot per digit bit (MATCH.ANY runs on the slow ADU pipe:
    // 77 % busy in the first ncu capture, profiles/r1_v1_radix_scatter.txt)
    u32 peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
    for (int b = 0; b < 8; b++) {
      bool bit = (d >> b) & 1u;
      u32 bal = __ballot_sync(0xffffffffu, bit);
      peers &= bit ? bal : ~bal;
    }
    // every lane reads its digit's running count (a broadcast for equal digits); the lowest peer bumps it
    const u32 below = peers & lt_mask;
    const u32 old = wcnt[w * 256 + d];
    if (valid && below == 0) wcnt[w * 256 + d] = old + (u32)__popc(peers);
    rank[r] = old + (u32)__popc(below);
    __syncwarp();
  }
  __syncthreads();
  {
    u32 run = 0;
    if (tid < 256) {
#pragma unroll
      for (int ww = 0; ww < RADIX_WARPS; ww++) {
        u32 c = wcnt[ww * 256 + tid];
        wcnt[ww * 256 + tid] = run;
        run += c;
      }
    }
    u32 tot;
    u32 ex = block_excl_scan(run, ws, tot);
    if (tid < 256) {
      texcl[tid] = ex;
      dbase[tid] = hbase - ex;
    }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < RADIX_ITEMS; r++) {
    u32 e = wbase + r * 32 + lane;
    if (e < N) {
      u32 d = (u32)(keys[r] >> shift) & mask;
      u32 lp = texcl[d] + wcnt[w * 256 + d] + rank[r];
      skeys[lp] = keys[r];
      svals[lp] = vin[e];
    }
  }
  __syncthreads();
  for (u32 j = tid; j < tile_n; j += RADIX_THREADS) {
    K key = skeys[j];
    u32 d = (u32)(key >> shift) & mask;
    u32 dst = dbase[d] + j;
    kout[dst] = key;
    vout[dst] = svals[j];
  }
}

// ---- persistent, double-buffered scatter.  The one-tile-per-block kernel above is latency bound (long-scoreboard
// stalls, 3 resident blocks per SM, profiles/r1_v2_radix_scatter.txt): loads, ranking and stores of a tile run one
// after the other.  Here each block walks tiles b, b+G, b+2G, ... and fetches the NEXT tile's keys and values
// with cp.async into the other half of a double buffer while it ranks and writes the current one, so every
// resident block always has a full tile (32-48 KB) in flight.
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(sa), "l"(gmem));
}
template <typename E>
__device__ __forceinline__ void tile_fetch(E* sdst, const E* __restrict__ g, u32 tile_base, u32 nvalid) {
  constexpr u32 PER = 16 / sizeof(E);
  const E* src = g + tile_base;
  for (u32 c = threadIdx.x; c < RADIX_TILE / PER; c += RADIX_THREADS) {
    u32 e0 = c * PER;
    if (e0 + PER <= nvalid) {
      cp_async16(sdst + e0, src + e0);
    } else {
      for (u32 t = 0; t < PER; t++)
        if (e0 + t < nvalid) sdst[e0 + t] = src[e0 + t];
    }
  }
}

// ---- TMA bulk copies (cp.async.bulk, SASS UBLKCP) for full tiles: ONE elected thread arms an mbarrier with the tile's byte
// count and issues two bulk copies (keys, values: 16 / 32 KB + 16 KB), the copy engine moves the tile and completes the
// barrier's transaction count, every thread waits on the barrier's phase — against 8-12 LDGSTS (cp.async 16 B) plus
// address arithmetic per thread and tile.  The shared-memory buffers are also written by ordinary stores (the staging
// of the previous tile): a fence.proxy.async orders those before the copy engine's writes.
__device__ __forceinline__ u32 smem_u32(const void* p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(u64* bar, u32 count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(u64* bar, u32 bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, u32 bytes, u64* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(u64* bar, u32 parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

template <typename K, int NMATCH, bool BULK = false>
__global__ void __launch_bounds__(RADIX_THREADS, (sizeof(K) == 4 ? RADIX_MINB32 : (RADIX_THREADS == 256 ? 2 : 1)))
k_radix_scatter_p(const K* __restrict__ kin, const u32* __restrict__ vin, K* __restrict__ kout, u32* __restrict__ vout,
                  u32 N, int shift, u32 mask, const u32* __restrict__ H, u32 ntiles) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  K* kbuf0 = reinterpret_cast<K*>(smem_raw);
  K* kbuf1 = kbuf0 + RADIX_TILE;
  u32* vbuf0 = reinterpret_cast<u32*>(kbuf1 + RADIX_TILE);
  u32* vbuf1 = vbuf0 + RADIX_TILE;
  u32* wcnt = vbuf1 + RADIX_TILE;          // [RADIX_WARPS][256]
  u32* texcl = wcnt + RADIX_WARPS * 256;   // [256]
  u32* dbase = texcl + 256;                // [256]
  u32* ws = dbase + 256;                   // [33]

  const u32 tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
  const u32 lt_mask = (1u << lane) - 1u;
  u32 tile = blockIdx.x;
  if (tile >= ntiles) return;
  __shared__ __align__(8) u64 s_bar[2];  // one per buffer (BULK)
  // fetch of one tile into buffer `b`: a full tile by two bulk copies, a ragged one by the per-thread cp.async loop
  auto fetch = [&](u32 t, u32 b) {
    const u32 tb = t * RADIX_TILE, nv = min((u32)RADIX_TILE, N - tb);
    K* kd = b ? kbuf1 : kbuf0;
    u32* vd = b ? vbuf1 : vbuf0;
    if (BULK && nv == RADIX_TILE) {
      if (tid == 0) {
        asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
        mbar_expect_tx(&s_bar[b], RADIX_TILE * (u32)(sizeof(K) + 4));
        bulk_g2s(kd, kin + tb, RADIX_TILE * (u32)sizeof(K), &s_bar[b]);
        bulk_g2s(vd, vin + tb, RADIX_TILE * 4u, &s_bar[b]);
      }
    } else {
      tile_fetch<K>(kd, kin, tb, nv);
      tile_fetch<u32>(vd, vin, tb, nv);
    }
    asm volatile("cp.async.commit_group;\n" ::);
  };
  u32 uses[2] = {0u, 0u};  // bulk fills waited for, per buffer (the barrier's phase parity)
  if (BULK) {
    if (tid == 0) {
      mbar_init(&s_bar[0], 1);
      mbar_init(&s_bar[1], 1);
      asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    }
    __syncthreads();
  }
  fetch(tile, 0);
  u32 hb_next = tid < 256 ? H[(u64)tid * ntiles + tile] : 0u;
  for (u32 it = 0; tile < ntiles; tile += gridDim.x, it++) {
    K* kb = (it & 1) ? kbuf1 : kbuf0;
    u32* vb = (it & 1) ? vbuf1 : vbuf0;
    const u32 hbase = hb_next;
    const u32 tile_base = tile * RADIX_TILE;
    const u32 tile_n = min((u32)RADIX_TILE, N - tile_base);
    asm volatile("cp.async.wait_group 0;\n" ::);
    if (BULK && tile_n == RADIX_TILE) {
      mbar_wait(&s_bar[it & 1], uses[it & 1] & 1u);
      uses[it & 1]++;
    }
    __syncthreads();  // the whole tile has landed
    K keys[RADIX_ITEMS];
    u32 vals[RADIX_ITEMS];
    u32 rank[RADIX_ITEMS];
    const u32 lbase = w * (32 * RADIX_ITEMS);
#pragma unroll
    for (int r = 0; r < RADIX_ITEMS; r++) {
      keys[r] = kb[lbase + r * 32 + lane];
      vals[r] = vb[lbase + r * 32 + lane];
    }
    for (u32 i = tid; i < RADIX_WARPS * 256; i += RADIX_THREADS) wcnt[i] = 0;
    __syncthreads();  // registers hold the tile: its buffer becomes the staging area
    // NMATCH < 0: peers are found through one shared-memory atomicOr per key on a [warp][digit] mask array instead of
    // eight ballots; the masks live in the tile's own key buffer, which is idle until the keys are staged
    u32* wmask = reinterpret_cast<u32*>(kb) + w * 256;
    if (NMATCH < 0) {
#pragma unroll
      for (int q = 0; q < 8; q++) wmask[q * 32 + lane] = 0;
      __syncwarp();
    }
    {
      u32 next = tile + gridDim.x;
      if (next < ntiles) {
        fetch(next, (it & 1) ^ 1u);
        if (tid < 256) hb_next = H[(u64)tid * ntiles + next];
      }
    }
#pragma unroll
    for (int r = 0; r < RADIX_ITEMS; r++) {
      bool valid = lbase + r * 32 + lane < tile_n;
      u32 d = (u32)(keys[r] >> shift) & mask;
      u32 peers;
      if (NMATCH < 0) {
        if (valid) atomicOr(&wmask[d], 1u << lane);
        __syncwarp();
        peers = valid ? wmask[d] : 0u;
        __syncwarp();
        if (valid && (peers & lt_mask) == 0) wmask[d] = 0;  // the lowest peer clears the mask for the next round
      } else if (r < NMATCH) {
        // a few rounds go through MATCH.ANY (ADU pipe) so that the ballot rounds (ALU pipe) are not the only load
        peers = __match_any_sync(0xffffffffu, valid ? d : 256u + lane);
      } else {
        peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
        for (int b = 0; b < 8; b++) {
          int m;  // 0 or -1: bit b of the digit, sign-extended (one bit-field extract)
          asm("bfe.s32 %0, %1, %2, 1;" : "=r"(m) : "r"(d), "r"(b));
          const u32 bal = __ballot_sync(0xffffffffu, m != 0);
          peers &= ~(bal ^ (u32)m);
        }
      }
      // every lane reads its digit's running count (a broadcast for equal digits); the lowest peer bumps it
      const u32 below = peers & lt_mask;
      const u32 old = wcnt[w * 256 + d];
      if (valid && below == 0) wcnt[w * 256 + d] = old + (u32)__popc(peers);
      rank[r] = old + (u32)__popc(below);
      __syncwarp();
    }
    __syncthreads();
    {
      u32 run = 0;
      if (tid < 256) {
#pragma unroll
        for (int ww = 0; ww < RADIX_WARPS; ww++) {
          u32 c = wcnt[ww * 256 + tid];
          wcnt[ww * 256 + tid] = run;
          run += c;
        }
      }
      u32 tot;
      u32 ex = block_excl_scan(run, ws, tot);
      if (tid < 256) {
        texcl[tid] = ex;
        dbase[tid] = hbase - ex;
      }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < RADIX_ITEMS; r++) {
      if (lbase + r * 32 + lane < tile_n) {
        u32 d = (u32)(keys[r] >> shift) & mask;
        u32 lp = texcl[d] + wcnt[w * 256 + d] + rank[r];
        kb[lp] = keys[r];
        vb[lp] = vals[r];
      }
    }
    __syncthreads();
    for (u32 j = tid; j < tile_n; j += RADIX_THREADS) {
      K key = kb[j];
      u32 d = (u32)(key >> shift) & mask;
      u32 dst = dbase[d] + j;
      kout[dst] = key;
      vout[dst] = vb[j];
    }
    // no barrier needed here: the next iteration only touches this buffer again after its own barriers
  }
}

template <typename K>
static size_t radix_smem_p() {  // (same for every NMATCH)
  return 2 * RADIX_TILE * (sizeof(K) + 4) + (RADIX_WARPS * 256 + 256 + 256 + 40) * 4;
}
static int g_radix_grid[2] = {0, 0};  // resident blocks of the persistent kernel (u32, u64 keys)

template <typename K>
static size_t radix_smem() {
  return RADIX_TILE * (sizeof(K) + 4) + (RADIX_WARPS * 256 + 256 + 256 + 40) * 4;
}

void radix_init() {
  CK(cudaFuncSetAttribute(k_radix_scatter<u32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)radix_smem<u32>()));
  CK(cudaFuncSetAttribute(k_radix_scatter<u64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)radix_smem<u64>()));
  CK(cudaFuncSetAttribute(k_radix_scatter_p<u32, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)radix_smem_p<u32>()));
  CK(cudaFuncSetAttribute(k_radix_scatter_p<u32, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)radix_smem_p<u32>()));
  CK(cudaFuncSetAttribute(k_radix_scatter_p<u32, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)radix_smem_p<u32>()));
  CK(cudaFuncSetAttribute(k_radix_scatter_p<u32, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)radix_smem_p<u32>()));
  CK(cudaFuncSetAttribute(k_radix_scatter_p<u64, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)radix_smem_p<u64>()));
  CK(cudaFuncSetAttribute(k_radix_scatter_p<u32, 4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)radix_smem_p<u32>()));
  CK(cudaFuncSetAttribute(k_radix_scatter_p<u64, 0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)radix_smem_p<u64>()));
  CK(cudaFuncSetAttribute(k_radix_scatter_p<u32, -1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)radix_smem_p<u32>()));
  CK(cudaFuncSetAttribute(k_radix_scatter_p<u64, -1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)radix_smem_p<u64>()));
  int dev = 0, sms = 0, b32 = 0, b64 = 0;
  CK(cudaGetDevice(&dev));
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b32, k_radix_scatter_p<u32, 0>, RADIX_THREADS, radix_smem_p<u32>()));
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b64, k_radix_scatter_p<u64, 0>, RADIX_THREADS, radix_smem_p<u64>()));
  g_radix_grid[0] = sms * (b32 > 0 ? b32 : 1);
  g_radix_grid[1] = sms * (b64 > 0 ? b64 : 1);
}

template <typename K>
static void radix_sort_impl(K*& keys, u32*& vals, K*& keys_alt, u32*& vals_alt, u32 N, int bit_lo, int bit_hi, u32* H,
                            u32* partials, cudaStream_t st) {
  if (N == 0) return;
  u32 ntiles = (u32)radix_tiles(N);
  for (int shift = bit_lo; shift < bit_hi; shift += 8) {
    int bits = bit_hi - shift < 8 ? bit_hi - shift : 8;
    u32 mask = (1u << bits) - 1u;
    {
      ProfScope ps(PK_RADIX_HIST, st);
      if (ps.p) ps.p->bytes[PK_RADIX_HIST] += sizeof(K) * (u64)N;
      k_radix_hist<K><<<ntiles, RADIX_THREADS, 0, st>>>(keys, N, shift, mask, H, ntiles);
      CK_LAUNCH();
    }
    scan_excl_u32(H, (u64)256 * ntiles, partials, nullptr, st);
    {
      ProfScope ps(PK_RADIX_SCATTER, st);
      if (ps.p) ps.p->bytes[PK_RADIX_SCATTER] += 2 * (sizeof(K) + 4) * (u64)N;  // read + write of (key, position)
      static const bool classic = getenv("KISS_B200_RADIX") && std::string(getenv("KISS_B200_RADIX")) == "classic";
      bool aligned = (((uintptr_t)keys | (uintptr_t)vals) & 15) == 0;
      if (classic || !aligned) {
        k_radix_scatter<K><<<ntiles, RADIX_THREADS, radix_smem<K>(), st>>>(keys, vals, keys_alt, vals_alt, N, shift,
                                                                          mask, H, ntiles);
      } else {
        u32 grid = (u32)g_radix_grid[sizeof(K) == 4 ? 0 : 1];
        if (grid > ntiles) grid = ntiles;
        static const int nmatch = [] { const char* e = getenv("KISS_B200_NMATCH"); return e ? atoi(e) : RADIX_ITEMS / 4; }();  // 0: 8.16 ms, 4: 7.94 ms, 8: 8.43 ms per 1 Gbp sort
        static const bool bulk = getenv("KISS_B200_RADIX_BULK") != nullptr;
        if (bulk && sizeof(K) == 4)
          k_radix_scatter_p<u32, 4, true><<<grid, RADIX_THREADS, radix_smem_p<K>(), st>>>((const u32*)keys, vals, (u32*)keys_alt, vals_alt, N, shift, mask, H, ntiles);
        else if (bulk)
          k_radix_scatter_p<K, 0, true><<<grid, RADIX_THREADS, radix_smem_p<K>(), st>>>(keys, vals, keys_alt, vals_alt, N, shift, mask, H, ntiles);
        else if (nmatch < 0)
          k_radix_scatter_p<K, -1><<<grid, RADIX_THREADS, radix_smem_p<K>(), st>>>(keys, vals, keys_alt, vals_alt, N, shift, mask, H, ntiles);
        else if (sizeof(K) == 4 && nmatch == 2)
          k_radix_scatter_p<u32, 2><<<grid, RADIX_THREADS, radix_smem_p<K>(), st>>>((const u32*)keys, vals, (u32*)keys_alt, vals_alt, N, shift, mask, H, ntiles);
        else if (sizeof(K) == 4 && nmatch == 4)
          k_radix_scatter_p<u32, 4><<<grid, RADIX_THREADS, radix_smem_p<K>(), st>>>((const u32*)keys, vals, (u32*)keys_alt, vals_alt, N, shift, mask, H, ntiles);
        else if (sizeof(K) == 4 && nmatch == 8)
          k_radix_scatter_p<u32, 8><<<grid, RADIX_THREADS, radix_smem_p<K>(), st>>>((const u32*)keys, vals, (u32*)keys_alt, vals_alt, N, shift, mask, H, ntiles);
        else
          k_radix_scatter_p<K, 0><<<grid, RADIX_THREADS, radix_smem_p<K>(), st>>>(keys, vals, keys_alt, vals_alt, N, shift,
                                                                                 mask, H, ntiles);
      }
      CK_LAUNCH();
    }
    g_launches += 2;
    K* tk = keys; keys = keys_alt; keys_alt = tk;
    u32* tv = vals; vals = vals_alt; vals_alt = tv;
  }
}

void radix_sort_u32(u32*& keys, u32*& vals, u32*& keys_alt, u32*& vals_alt, u32 N, int bit_lo, int bit_hi, u32* H,
                    u32* partials, cudaStream_t st) {
  radix_sort_impl<u32>(keys, vals, keys_alt, vals_alt, N, bit_lo, bit_hi, H, partials, st);
}
void radix_sort_u64(u64*& keys, u32*& vals, u64*& keys_alt, u32*& vals_alt, u32 N, int bit_lo, int bit_hi, u32* H,
                    u32* partials, cudaStream_t st) {
  radix_sort_impl<u64>(keys, vals, keys_alt, vals_alt, N, bit_lo, bit_hi, H, partials, st);
}

}  // namespace kiss
