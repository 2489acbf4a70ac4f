// lms.cu — K2: L/S-type classification and LMS extraction on the packed text.
//
// Replaces get_lms (kiss_common.hpp:483-541 per-thread scan, 543-579 driver).  The reference walks the
// text right to left with a 2-bit type shift register (:516); here the same recurrence
//     type(i) = S  iff  S[i] < S[i+1]  or  (S[i] == S[i+1] and type(i+1) = S),   type(n-1) = L
// is a carry chain: per base "generate" = (S[i] < S[i+1]), "propagate" = (S[i] == S[i+1]), so the S-type
// bits of 32 bases are the carries of ONE 64-bit integer addition (bases sit at descending bit
// positions, so text-right-to-left is the adder's low-to-high carry direction).  Carries between words
// and between 2048-base chunks are resolved by a (generate, propagate) summary per chunk — the
// chunk-boundary look-ahead of the reference (kiss_common.hpp:502-504) in associative form.
// Position i is LMS iff S[i-1] > S[i] and type(i) = S (:518-522).
//
// Outputs: one mask word per text word (LMS bits), LMS count per chunk, count[c] (bucket[4*256+c],
// :529-536) and lms_count[c] (bucket[LMS_TYPE*256+c], :519).
// Algorithmic bytes: n/4 read per pass over the packed text; 4 m + 4 m written by the emit pass.
#include "common.cuh"
#include "kernels.h"

namespace kiss {

// S-type bits of word j given the carry from the word to its right; returns the carry out (type of the
// word's first base).  `next_top` = first base of word j+1.
__device__ __forceinline__ u32 word_types(u64 x, u64 next_top, u64 j, u64 n, u32 cin, u64& s_bits, bool& all_prop) {
  u64 y = (x << 2) | next_top;
  u64 lt, eq;
  field_cmp(x, y, lt, eq);
  // bases with a successor inside the text: i < n - 1
  u64 first = j * 32;
  u32 cntv = (n - 1 > first) ? (u32)min((u64)32, n - 1 - first) : 0u;
  u64 vm = first_fields(cntv);
  lt &= vm;
  eq &= vm;
  all_prop = (eq == FIELD_LO);
  u64 a = (lt | eq) * 3ull, b = lt;
  u64 s1 = a + b;
  u32 c1 = s1 < a;
  u64 s2 = s1 + cin;
  u32 c2 = s2 < s1;
  u32 cout = c1 | c2;
  u64 carries = s2 ^ a ^ b;  // carry INTO every bit
  s_bits = ((carries >> 2) & FIELD_LO) | ((u64)cout << 62);
  return cout;
}

// pass A: (generate, propagate) of every chunk
__global__ void __launch_bounds__(256) k_lms_summary(const u64* __restrict__ T, u64 n, u64 nwords, u64 nchunks,
                                                    u8* __restrict__ gp) {
  u64 c = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= nchunks) return;
  u64 w0 = c * LMS_CHUNK_WORDS, w1 = min(nwords, w0 + LMS_CHUNK_WORDS);
  u32 carry = 0;
  bool prop = true;
  u64 nxt = T[w1];  // padded: always readable
  for (u64 j = w1; j-- > w0;) {
    u64 x = T[j];
    u64 s;
    bool ap;
    carry = word_types(x, nxt >> 62, j, n, carry, s, ap);
    prop = prop && ap;
    nxt = x;
  }
  gp[c] = (u8)(carry | (prop ? 2u : 0u));
}

// pass B: carry into every chunk (bit 2 of gp[c]); single block
__global__ void __launch_bounds__(1024) k_lms_carry(u8* __restrict__ gp, u64 nchunks) {
  __shared__ u8 rg[1024], rp[1024], rc[1024];
  u64 per = (nchunks + 1023) / 1024;
  u64 c0 = min(nchunks, (u64)threadIdx.x * per), c1 = min(nchunks, c0 + per);
  u32 g = 0, p = 1;
  for (u64 c = c1; c-- > c0;) {
    u32 v = gp[c];
    g = (v & 1u) | (((v >> 1) & 1u) & g);
    p = p & ((v >> 1) & 1u);
  }
  rg[threadIdx.x] = (u8)g;
  rp[threadIdx.x] = (u8)p;
  __syncthreads();
  if (threadIdx.x == 0) {
    u32 carry = 0;
    for (int t = 1023; t >= 0; t--) {
      rc[t] = (u8)carry;
      carry = rg[t] | (rp[t] & carry);
    }
  }
  __syncthreads();
  u32 carry = rc[threadIdx.x];
  for (u64 c = c1; c-- > c0;) {
    u32 v = gp[c];
    gp[c] = (u8)(v | (carry << 2));
    carry = (v & 1u) | (((v >> 1) & 1u) & carry);
  }
}

// pass C: LMS masks, per-chunk LMS counts, symbol totals
__global__ void __launch_bounds__(256) k_lms_mask(const u64* __restrict__ T, u64 n, u64 nwords, u64 nchunks,
                                                 const u8* __restrict__ gp, u64* __restrict__ lmsmask,
                                                 u32* __restrict__ chunk_cnt, u64* __restrict__ counters) {
  u64 c = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  u32 cnt[4] = {0, 0, 0, 0}, lcnt[4] = {0, 0, 0, 0};
  if (c < nchunks) {
    u64 w0 = c * LMS_CHUNK_WORDS, w1 = min(nwords, w0 + LMS_CHUNK_WORDS);
    u32 carry = (gp[c] >> 2) & 1u;
    u64 nxt = T[w1];
    u32 total = 0;
    for (u64 j = w1; j-- > w0;) {
      u64 x = T[j];
      u64 s;
      bool ap;
      carry = word_types(x, nxt >> 62, j, n, carry, s, ap);
      nxt = x;
      // S[i-1] aligned under S[i]; word 0 has no predecessor: repeat its own first base (never greater)
      u64 prev_last = (j > 0) ? (T[j - 1] & 3ull) : (x >> 62);
      u64 z = (x >> 2) | (prev_last << 62);
      u64 ltp, eqp;
      field_cmp(x, z, ltp, eqp);  // S[i] < S[i-1]
      u64 lms = ltp & s;
      lmsmask[j] = lms;
      total += (u32)__popcll(lms);
      u64 first = j * 32;
      u32 cn = (n > first) ? (u32)min((u64)32, n - first) : 0u;
      u64 vn = first_fields(cn);
      u64 xh = (x >> 1) & FIELD_LO, xl = x & FIELD_LO;
      u64 m3 = xh & xl, m2 = xh & ~xl, m1 = ~xh & xl, m0 = FIELD_LO & ~(xh | xl);
      cnt[0] += (u32)__popcll(m0 & vn); cnt[1] += (u32)__popcll(m1 & vn);
      cnt[2] += (u32)__popcll(m2 & vn); cnt[3] += (u32)__popcll(m3 & vn);
      lcnt[0] += (u32)__popcll(m0 & lms); lcnt[1] += (u32)__popcll(m1 & lms);
      lcnt[2] += (u32)__popcll(m2 & lms); lcnt[3] += (u32)__popcll(m3 & lms);
    }
    chunk_cnt[c] = total;
  }
  // warp reduce, one atomic per warp and counter
#pragma unroll
  for (int s = 0; s < 4; s++) {
    u32 a = cnt[s], b = lcnt[s];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      a += __shfl_xor_sync(0xffffffffu, a, o);
      b += __shfl_xor_sync(0xffffffffu, b, o);
    }
    if (lane_id() == 0) {
      if (a) atomicAdd(&counters[s], (u64)a);
      if (b) atomicAdd(&counters[4 + s], (u64)b);
    }
  }
}

void lms_classify(const u64* T, u64 n, u8* gp, u64* lmsmask, u32* chunk_cnt, u64* counters, cudaStream_t st) {
  u64 nwords = (n + 31) / 32, nchunks = lms_chunks(n);
  if (nchunks == 0) return;
  ProfScope ps(PK_LMS_CLASSIFY, st);
  unsigned grid = (unsigned)((nchunks + 255) / 256);
  k_lms_summary<<<grid, 256, 0, st>>>(T, n, nwords, nchunks, gp);
  CK_LAUNCH();
  k_lms_carry<<<1, 1024, 0, st>>>(gp, nchunks);
  CK_LAUNCH();
  k_lms_mask<<<grid, 256, 0, st>>>(T, n, nwords, nchunks, gp, lmsmask, chunk_cnt, counters);
  CK_LAUNCH();
  g_launches += 3;
}

// pass E: positions (ascending) and first-16-base keys.  One block = 256 mask words (4 chunks); the positions
// of the tile are staged in shared memory and written with coalesced stores.
#define EMIT_WORDS 256
__global__ void __launch_bounds__(EMIT_WORDS) k_lms_emit(const u64* __restrict__ T, u64 nwords,
                                                        const u64* __restrict__ lmsmask,
                                                        const u32* __restrict__ chunk_off, u32* __restrict__ pos,
                                                        u32* __restrict__ key) {
  __shared__ u32 ws[33];
  __shared__ u32 spos[EMIT_WORDS * 16], skey[EMIT_WORDS * 16];  // at most 16 LMS positions per 32 bases
  const u64 j = (u64)blockIdx.x * EMIT_WORDS + threadIdx.x;
  u64 mk = j < nwords ? lmsmask[j] : 0ull;
  u32 total;
  u32 ex = block_excl_scan((u32)__popcll(mk), ws, total);
  if (mk) {
    u64 a = T[j], b = T[j + 1];
    while (mk) {
      int lz = __clzll((long long)mk);
      u32 f = (u32)lz >> 1;  // field index = base offset in the word
      mk &= ~(1ull << (63 - lz));
      u32 sh = 2 * f;
      u64 wv = sh ? ((a << sh) | (b >> (64 - sh))) : a;
      spos[ex] = (u32)(j * 32 + f);
      skey[ex] = (u32)(wv >> 32);
      ex++;
    }
  }
  __syncthreads();
  const u32 off = chunk_off[(u64)blockIdx.x * (EMIT_WORDS / LMS_CHUNK_WORDS)];
  for (u32 t = threadIdx.x; t < total; t += EMIT_WORDS) {
    pos[off + t] = spos[t];
    key[off + t] = skey[t];
  }
}

void lms_emit(const u64* T, u64 n, const u64* lmsmask, const u32* chunk_off, u32* pos, u32* key, cudaStream_t st) {
  u64 nwords = (n + 31) / 32;
  if (nwords == 0) return;
  ProfScope ps(PK_LMS_EMIT, st);
  k_lms_emit<<<(unsigned)((nwords + EMIT_WORDS - 1) / EMIT_WORDS), EMIT_WORDS, 0, st>>>(T, nwords, lmsmask, chunk_off,
                                                                                       pos, key);
  CK_LAUNCH();
  g_launches++;
}

__global__ void k_count_le(const u32* __restrict__ pos, u32 cnt, u32 limit, u32* __restrict__ out) {
  u32 lo = 0, hi = cnt;  // first index with pos > limit
  while (lo < hi) {
    u32 mid = lo + (hi - lo) / 2;
    if (pos[mid] <= limit) lo = mid + 1; else hi = mid;
  }
  *out = lo;
}

void count_le(const u32* pos, u32 cnt, u32 limit, u32* out, cudaStream_t st) {
  k_count_le<<<1, 1, 0, st>>>(pos, cnt, limit, out);
  CK_LAUNCH();
  g_launches++;
}

}  // namespace kiss
