// induce.cu — K4 placement and K5 induction.
//
// Replaces put_lms_suffix (kiss_common.hpp:445-481) and induced_sort (kiss_common.hpp:422-442: induced_L
// 372-403 / 226-246, induced_clear 405-420, induced_S 192-224 / 40-61).
//
// The reference scans SA slot by slot and, for a filled slot v, appends v-1 to the cursor of its bucket when
// the type test holds (L pass: S[v-1] >= S[v], :241; S pass: S[v-1] <= S[v], :55-56).  A maximal run of filled
// slots can be processed as one *wave* (the reference's own block scheme, :389-393 / :209-213): everything a
// wave appends lands on slots that were empty when the wave was cut.  Here a wave is a stable 4-way multisplit
// on the GPU: per 2048-entry tile count the induced suffixes per symbol, scan the tile counts, scatter with
// ballot ranks.  Wave boundaries are the bucket cursors themselves, so no EMPTY markers, no induced_clear and
// none of the reference's artefact EMPTY writes (:241) are needed; the S pass overwrites the LMS slots.
// SURVEY.md A.4 holds the equivalence argument; tests/model_gpu_algo.py is the same schedule in numpy.
//
// Algorithmic bytes per scanned entry: 4 (SA read) + 0.25 (packed text) + 4 per induced suffix; the context
// array adds 4 + 4 of streaming traffic and removes all but ~0.3 n of the random text gathers.
#include <cstdio>
#include <cstdlib>
#include "common.cuh"
#include "kernels.h"

namespace kiss {

// ------------------------------------------------------------------------------------------- placement
__device__ __forceinline__ u64 lms_slot(const Buckets& bk, u64 r) {
  // r = rank among the sorted LMS suffixes (sentinel excluded); they are grouped by first symbol.
  // (selects, not indexing: a dynamically indexed kernel parameter is spilled to local memory)
  u64 end = bk.B[1], cnt = bk.lms_count[0], off = bk.lms_off[0];
  if (r >= bk.lms_off[1]) { end = bk.B[2]; cnt = bk.lms_count[1]; off = bk.lms_off[1]; }
  if (r >= bk.lms_off[2]) { end = bk.B[3]; cnt = bk.lms_count[2]; off = bk.lms_off[2]; }
  if (r >= bk.lms_off[3]) { end = bk.B[4]; cnt = bk.lms_count[3]; off = bk.lms_off[3]; }
  return end - cnt + (r - off);
}

#define PLACE_INS_SMEM 2048u
__global__ void __launch_bounds__(256) k_place_lms(const u32* __restrict__ longs, u32 nl,
                                                  const u32* __restrict__ sorted_shorts, const u32* __restrict__ ins,
                                                  u32 ns, Buckets bk, u64 n, u32* __restrict__ SA,
                                                  u32* __restrict__ CTX) {
  __shared__ u32 sins[PLACE_INS_SMEM];
  const bool cached = ns <= PLACE_INS_SMEM;
  if (cached)
    for (u32 t = threadIdx.x; t < ns; t += blockDim.x) sins[t] = ins[t];
  __syncthreads();
  const u32* I = cached ? sins : ins;
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0) {
    SA[0] = (u32)n;
    if (CTX) CTX[0] = 1u;
  }
  u64 slot;
  u32 val;
  if (i < nl) {
    // shorts emitted before long i: those with ins <= i  (ins is non-decreasing)
    u32 lo = 0, hi = ns;
    if (ns && (u32)i < I[0]) hi = 0;
    else if (ns && (u32)i >= I[ns - 1]) lo = ns;
    while (lo < hi) {
      u32 mid = lo + (hi - lo) / 2;
      if (I[mid] <= (u32)i) lo = mid + 1; else hi = mid;
    }
    slot = lms_slot(bk, i + lo);
    val = longs[i];
  } else if (i < (u64)nl + ns) {
    u32 s = (u32)(i - nl);
    slot = lms_slot(bk, (u64)I[s] + s);
    val = sorted_shorts[s];
  } else {
    return;
  }
  SA[slot] = val;
  if (CTX) CTX[slot] = 1u;  // context not fetched yet (see the waves below)
}

void place_lms(const u32* longs, u32 nl, const u32* sorted_shorts, const u32* ins, u32 ns, Buckets bk, u64 n, u32* SA,
               u32* CTX, cudaStream_t st) {
  u64 tot = (u64)nl + ns;
  if (tot == 0) tot = 1;
  ProfScope ps(PK_PLACE, st);
  k_place_lms<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(longs, nl, sorted_shorts, ins, ns, bk, n, SA, CTX);
  CK_LAUNCH(); g_launches++;
}

void place_sorted(const u32* sorted, u32 cnt, Buckets bk, u64 n, u32* SA, u32* CTX, cudaStream_t st) {
  place_lms(sorted, cnt, nullptr, nullptr, 0, bk, n, SA, CTX, st);
}

// ------------------------------------------------------------------------------------------- waves
// Every SA entry v travels with a context word CTX = 1 << 2L | (S[v-L] .. S[v-1]) holding L <= 15 bases
// that precede v (S[v-1] in the low two bits).  Scanning v needs only S[v-1] (the bucket gives S[v]), and the
// suffix v-1 it induces inherits CTX >> 2 — so the packed text is gathered (one random 32-byte sector) only
// when a context runs out (CTX == 1): once per LMS suffix and once per 15 steps of a chain, instead of once
// per scanned entry as in the reference's prepare loops (kiss_common.hpp:63-113, 278-325).

#define WAVE_THREADS (WAVE_TILE / 16)
#define WAVE_ITEMS 16
#define WAVE_WARPS (WAVE_THREADS / 32)
#define WAVE_MINB (WAVE_TILE == 4096 ? 4 : WAVE_TILE == 8192 ? 2 : 1)

struct WaveArgs {
  u32* SA;
  u32* CTX;
  const u32* inSA;   // scanned entries: SA itself, or the sorted LMS list (then lo/len index that list)
  const u32* inCTX;  // nullptr: every scanned entry fetches its context (LMS list)
  const u64* T;
  u64 lo;       // first slot of the wave
  u64 len;      // number of slots
  u32 down;     // scan from the top slot downwards
  u32 symmask;  // symbols that may be induced
  u64 cur[4];   // bucket cursors
  // the scanned list as the sorted long suffixes with the few short ones inserted on the fly (VIRT instances of the
  // wave kernel): slot L holds shorts[t] when ins[t] + t == L, else inSA[L - (number of shorts placed before L)]
  const u32* shorts;
  const u32* ins;
  u32 ns;
};

// context of suffix v: the 15 bases before it (two adjacent text words).  Variants that were measured and lost:
// a one-word context (what the word holding S[v-1] offers: short contexts run out in the middle of chains, every
// exhausted entry costs another random access later, 45 % slower overall); a copy of the text in overlapping
// 32-base windows so that one aligned load suffices (a DRAM miss fetches ~128 B whatever the request size —
// 107 B per gather against 93 B here, and the copy is twice the text so fewer gathers hit L2);
// branch-free batched gathers (no gain: the LMS waves are bound by random DRAM accesses, not by latency);
// fetching the LMS contexts up front in 2-8 sweeps, each restricted to an L2-sized slice of the text
// (2 sweeps: 3.8 ms + 1.7 ms of gather-free LMS waves against 5.9 ms with the gathers in the waves — a wash).
// (L2 fetch-size hints on the gather loads — ld.global.nc.L2::64B / 128B / 256B — measured no different)
__device__ __forceinline__ u64 ldg_gather(const u64* p) { return __ldg(p); }
// the 32 bases from p on, of which the caller uses the first 15: the second word is fetched only when they
// straddle the word boundary
__device__ __forceinline__ u64 word_at_gather(const u64* __restrict__ T, u64 p) {
  u64 i = p >> 5;
  u32 s = (u32)(p & 31) * 2;
  u64 a = ldg_gather(T + i);
  if (s <= 34) return a << s;
  u64 b = ldg_gather(T + i + 1);
  return (a << s) | (b >> (64 - s));
}
__device__ __forceinline__ u32 ctx_gather(const u64* __restrict__ T, u32 v) {
  if (v >= 15) return (1u << 30) | (u32)(word_at_gather(T, (u64)v - 15) >> 34);
  if (v == 0) return 1u;
  return (1u << (2 * v)) | (u32)(__ldg(T) >> (64 - 2 * v));
}

// symbol (0..3) the entry induces into, or 4
__device__ __forceinline__ u32 ctx_symbol(u32 c, u32 symmask) {
  u32 cu = c & 3u;
  return (c > 1u && ((symmask >> cu) & 1u)) ? cu : 4u;
}

// ---- runs of the bucket's own symbol (the on-device tail, below)
// Inside a bucket a wave's successor is made of the entries whose preceding base is the bucket's symbol, so a run of r
// such bases before a suffix is a chain of r waves — a 10^6-base homopolymer run (or a run of N, which the reference
// maps to A) is 10^6 waves of one entry each.  When EVERY entry of a wave has at least K such bases before it, the next
// K waves are known without scanning them: wave k holds the suffixes v_j - k in the order of the current wave, nothing
// goes to another bucket.  They are written in one burst and the tail continues at wave K.
// bases of the bucket's symbol right before suffix v, as far as its context word shows (exact up to the context length)
__device__ __forceinline__ u32 ctx_len(u32 c) { return c > 1u ? (31u - (u32)__clz(c)) >> 1 : 0u; }
__device__ __forceinline__ u32 ctx_run(u32 c, u32 sym) {
  if (c <= 1u) return 0u;
  const u32 x = c ^ (sym * 0x55555555u);
  const u32 tz = x ? ((u32)(__ffs((int)x) - 1) >> 1) : 16u;
  return min(tz, ctx_len(c));
}
// the same from the text: the exact run when it is shorter than cap, else a lower bound in [cap, cap + 32)
#define RUN_CAP 4096u
#define BURST_MIN 3u
__device__ __noinline__ u32 text_run(const u64* __restrict__ T, u32 v, u32 sym, u32 cap) {
  const u64 pat = (u64)sym * 0x5555555555555555ull;
  u32 cnt = 0;
  u64 pos = v;
  while (cnt < cap && pos >= 32) {
    const u64 x = word_at(T, pos - 32) ^ pat;  // bases pos-32 .. pos-1, the last one in the lowest bits
    const u32 tz = x ? ((u32)(__ffsll((long long)x) - 1) >> 1) : 32u;
    cnt += tz;
    pos -= tz;
    if (tz < 32) return cnt;
  }
  while (cnt < cap && pos > 0 && base_at(T, pos - 1) == sym) { cnt++; pos--; }  // (the head of the text)
  return cnt;
}
// lower bound of the run before suffix v: what the previous wave handed over, what the context shows, and — when the
// context is all the bucket's symbol and nothing was handed over — what the text shows
__device__ __forceinline__ u32 run_bound(const u64* __restrict__ T, u32 v, u32 c, u32 handed, u32 sym) {
  const u32 cr = ctx_run(c, sym);
  u32 rr = max(handed, cr);
  if (handed == 0 && cr != 0 && cr == ctx_len(c)) rr = max(rr, text_run(T, v, sym, RUN_CAP));
  return rr;
}

struct WaveSmem {
  u32 sb[5];
  u32 sSA[WAVE_TILE];
  u32 sCTX[WAVE_TILE];
};


// ---- single-pass wave: count, decoupled look-back across tiles, scatter — one read of (CTX, SA) per entry.
// Tiles take their index from an atomic ticket (so a tile never waits on one that has not started), publish
// their per-symbol counts in a 64-bit status word [epoch:30 | flag:2 | count:32] and look back over their
// predecessors' words for the exclusive prefix.  The epoch (one per wave) makes stale words of earlier waves
// read as "empty", so the status array is zeroed once per call, not per wave.
#define WF_LB 1  // look-back window = 32 * WF_LB tiles (4: every lane polling 4 predecessors measured 15 % slower)
#define WF_AGG 1ull
#define WF_INCL 2ull
__device__ __forceinline__ u64 wf_pack(u32 epoch, u64 flag, u32 v) { return ((u64)epoch << 34) | (flag << 32) | v; }

// Thread t of a tile owns 16 CONSECUTIVE slots (four 16-byte loads per array), so ranking needs no warp votes:
// a packed 4 x 16-bit counter per thread, one warp scan and one cross-warp prefix give every induced suffix its
// place in the tile (the first version spent 68 % of its issue slots on ballots and popcounts,
// profiles/r1_v3_wave_fused.txt).  Tiles are cut on 4-slot boundaries of the absolute slot index; the slots
// outside [lo, lo+len) at either end are masked.
struct WaveShared {
  WaveSmem sm;
  u64 wsum[WAVE_WARPS];
  u32 off[4];
  u32 tile;
  // on-device tail loop
  u64 cur[4];
  u64 p;
  u32 nwaves;
  u32 ntile, nmid;  // statistics: waves that took the tile path / the one-entry-per-thread path
  u32 wmin[WAVE_WARPS];  // per-warp minimum of the run bounds (burst decision of the one-entry-per-thread path)
};

// one tile of a wave.  LOOKBACK = false: the wave is this single tile (no status traffic); tot[] gets the counts.
template <bool DOWN, bool LOOKBACK, bool VIRT = false>
__device__ __forceinline__ void wave_tile(const WaveArgs& a, const u32 tile, WaveShared& sh, u64* __restrict__ status,
                                          u32* __restrict__ state, u32 ntiles, u32 epoch, u32* __restrict__ tot) {
  WaveSmem& sm = sh.sm;
  u64* s_wsum = sh.wsum;
  u32* s_off = sh.off;
  const u32 w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const u64 hi = a.lo + a.len;
  // logical index q of the aligned wave: up: slot = abase + q; down: slot = atop - 1 - q
  // (cut on 16-byte boundaries of the scanned array: `mis` = its misalignment in entries, 0 for SA itself)
  const u64 mis = VIRT ? 0ull : ((u64)(uintptr_t)a.inSA >> 2) & 3ull;
  const u64 abase = a.lo - ((a.lo + mis) & 3ull), atop = hi + ((4ull - ((hi + mis) & 3ull)) & 3ull);
  const u64 q0 = (u64)tile * WAVE_TILE + (u64)threadIdx.x * WAVE_ITEMS;
  u32 cx[WAVE_ITEMS], val[WAVE_ITEMS];
  if constexpr (VIRT) {
    // (upward only) the thread's 16 logical slots of the merged list; t = short suffixes placed before the slot
    typedef long long s64;
    const s64 first = (s64)abase + (s64)q0, slo = (s64)a.lo, shi = (s64)hi;
    const s64 L0 = first > slo ? first : slo;
    u32 t = 0;
    {
      u32 lo_ = 0, hi_ = a.ns;  // first short whose slot ins[s] + s is >= L0
      while (lo_ < hi_) {
        const u32 mid = lo_ + (hi_ - lo_) / 2;
        if ((s64)a.ins[mid] + mid < L0) lo_ = mid + 1; else hi_ = mid;
      }
      t = lo_;
    }
#pragma unroll
    for (int i = 0; i < WAVE_ITEMS; i++) {
      const s64 L = first + i;
      u32 c = 1u, v = 0u;
      if (L >= slo && L < shi) {
        if (t < a.ns && (s64)a.ins[t] + t == L) {
          v = a.shorts[t];
          t++;
        } else {
          v = a.inSA[L - t];
          if (a.inCTX) c = a.inCTX[L - t];
        }
      }
      cx[i] = c;
      val[i] = v;
    }
  } else {
    // address-ascending block of 16 slots owned by this thread (signed: the last tile of a downward wave may
    // reach below slot 0)
    typedef long long s64;
    const s64 first = DOWN ? (s64)atop - (s64)q0 - WAVE_ITEMS : (s64)abase + (s64)q0;
    const s64 slo = (s64)a.lo, shi = (s64)hi;
#pragma unroll
    for (int g = 0; g < WAVE_ITEMS / 4; g++) {
      u32 c4[4] = {1u, 1u, 1u, 1u}, v4[4] = {0u, 0u, 0u, 0u};
      const s64 s0 = first + 4 * g;
      if (s0 >= slo && s0 + 4 <= shi) {
        uint4 v = *reinterpret_cast<const uint4*>(a.inSA + s0);
        v4[0] = v.x; v4[1] = v.y; v4[2] = v.z; v4[3] = v.w;
        if (a.inCTX) {
          uint4 c = *reinterpret_cast<const uint4*>(a.inCTX + s0);
          c4[0] = c.x; c4[1] = c.y; c4[2] = c.z; c4[3] = c.w;
        }
      } else if (s0 + 4 > slo && s0 < shi) {
#pragma unroll
        for (int e = 0; e < 4; e++) {
          s64 sl = s0 + e;
          if (sl >= slo && sl < shi) { c4[e] = a.inCTX ? a.inCTX[sl] : 1u; v4[e] = a.inSA[sl]; }
        }
      }
#pragma unroll
      for (int e = 0; e < 4; e++) {
        const int i = DOWN ? (WAVE_ITEMS - 1 - (4 * g + e)) : (4 * g + e);  // scan order inside the thread
        cx[i] = c4[e];
        val[i] = v4[e];
      }
    }
  }
  // symbols + per-thread counts: 8-bit fields while counting (<= 16 per thread), 16-bit fields for the scans
  u64 symbits = 0;
  u32 P8 = 0;
#pragma unroll
  for (int i = 0; i < WAVE_ITEMS; i++) {
    u32 c = cx[i], v = val[i];
    if (c == 1u && v != 0 && v != KISS_EMPTY) {  // context ran out (entries outside the wave have v == 0)
      c = ctx_gather(a.T, v);
      cx[i] = c;
    }
    u32 cu = ctx_symbol(c, a.symmask);
    symbits |= (u64)cu << (4 * i);
    P8 += (cu < 4) ? (1u << (8 * cu)) : 0u;
  }
  const u64 P = (u64)(P8 & 0xFFu) | ((u64)(P8 & 0xFF00u) << 8) | ((u64)(P8 & 0xFF0000u) << 16) | ((u64)(P8 >> 24) << 48);
  // exclusive scan of P over the threads of the warp, warp totals to shared memory
  u64 incl = P;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    u64 t = __shfl_up_sync(0xffffffffu, incl, o);
    if ((int)lane >= o) incl += t;
  }
  if (lane == 31) s_wsum[w] = incl;
  __syncthreads();
  if (threadIdx.x == 0) {
    u64 totp = 0;
    for (int ww = 0; ww < WAVE_WARPS; ww++) totp += s_wsum[ww];  // fields <= 4096: no carry between fields
    u32 acc = 0;
    for (int s = 0; s < 4; s++) {
      sm.sb[s] = acc;
      acc += (u32)(totp >> (16 * s)) & 0xFFFFu;
    }
    sm.sb[4] = acc;
  }
  __syncthreads();
  // ---- look-back: warp s (< 4) resolves symbol s
  if (!LOOKBACK) {
    if (threadIdx.x < 4) {
      s_off[threadIdx.x] = 0;
      tot[threadIdx.x] = sm.sb[threadIdx.x + 1] - sm.sb[threadIdx.x];
    }
  } else if (w < 4) {
    const u32 s = w;
    const u32 mine = sm.sb[s + 1] - sm.sb[s];
    volatile u64* st = status;
    if (lane == 0) st[(u64)tile * 4 + s] = wf_pack(epoch, tile == 0 ? WF_INCL : WF_AGG, mine);
    u32 excl = 0;
    if (tile > 0) {
      long long base = (long long)tile - 1;
      u32 spins = 0;
      for (;;) {
        u64 word[WF_LB];
        bool ready;
        do {
          ready = true;
#pragma unroll
          for (int k = 0; k < WF_LB; k++) {
            const long long t = base - lane - 32 * k;
            word[k] = (t >= 0) ? st[(u64)t * 4 + s] : wf_pack(epoch, WF_INCL, 0);
            ready &= (u32)(word[k] >> 34) == epoch && ((word[k] >> 32) & 3ull) != 0;
          }
          if (++spins > (1u << 24)) {  // watchdog: never hang
            state[1] = 1u;
            tot[4] = 1u;  // (pinned host memory, read after the last wave)
            ready = true;
#pragma unroll
            for (int k = 0; k < WF_LB; k++) word[k] = wf_pack(epoch, WF_INCL, 0);
          }
        } while (!__all_sync(0xffffffffu, ready));
        bool done = false;
#pragma unroll
        for (int k = 0; k < WF_LB; k++) {
          if (!done) {  // (warp-uniform)
            const u32 inclm = __ballot_sync(0xffffffffu, ((word[k] >> 32) & 3ull) == WF_INCL);
            u32 v = (u32)word[k];
            if (inclm) {
              const u32 first = __ffs(inclm) - 1;  // nearest predecessor holding an inclusive prefix
              if (lane > first) v = 0;
              done = true;
            }
            excl += __reduce_add_sync(0xffffffffu, v);
          }
        }
        if (done) break;
        base -= 32 * WF_LB;
      }
      if (lane == 0) st[(u64)tile * 4 + s] = wf_pack(epoch, WF_INCL, excl + mine);
    }
    if (lane == 0) {
      s_off[s] = excl;
      if (tile == ntiles - 1) tot[s] = excl + mine;
    }
  }
  // tile-local position of this thread's first induced suffix per symbol: symbol base + earlier warps + earlier
  // threads of the warp; inside the thread an 8-bit running count per symbol
  u64 R = incl - P;
  for (u32 ww = 0; ww < w; ww++) R += s_wsum[ww];
  const u32 r0 = ((u32)R & 0xFFFFu) + sm.sb[0], r1 = ((u32)(R >> 16) & 0xFFFFu) + sm.sb[1],
            r2 = ((u32)(R >> 32) & 0xFFFFu) + sm.sb[2], r3 = (u32)(R >> 48) + sm.sb[3];
  u32 L8 = 0;
#pragma unroll
  for (int i = 0; i < WAVE_ITEMS; i++) {
    u32 cu = (u32)(symbits >> (4 * i)) & 7u;
    if (cu < 4) {
      u32 lp = (cu == 0 ? r0 : cu == 1 ? r1 : cu == 2 ? r2 : r3) + ((L8 >> (8 * cu)) & 0xFFu);
      L8 += 1u << (8 * cu);
      sm.sSA[lp] = val[i] - 1u;
      sm.sCTX[lp] = cx[i] >> 2;
    }
  }
  __syncthreads();
  // each symbol's run goes out with coalesced stores
#pragma unroll
  for (int s = 0; s < 4; s++) {
    const u32 b = sm.sb[s], e = sm.sb[s + 1];
    const long long cur = (long long)a.cur[s], off = (long long)s_off[s] - (long long)b;
    const long long base = DOWN ? cur - 1 - off : cur + off;  // dst = base -/+ idx
    for (u32 idx = b + threadIdx.x; idx < e; idx += WAVE_THREADS) {
      const long long dst = DOWN ? base - (long long)idx : base + (long long)idx;
      a.SA[dst] = sm.sSA[idx];
      a.CTX[dst] = sm.sCTX[idx];
    }
  }
}

template <bool DOWN, int MINB, bool VIRT = false>
__global__ void __launch_bounds__(WAVE_THREADS, MINB) k_wave_fused(WaveArgs a, u64* __restrict__ status, u32* __restrict__ state,
                                                                  u32 tile_base, u32 ntiles, u32 epoch, u32* __restrict__ tot) {
  extern __shared__ __align__(16) unsigned char wave_smem[];
  WaveShared& sh = *reinterpret_cast<WaveShared*>(wave_smem);
  if (threadIdx.x == 0) sh.tile = atomicAdd(&state[0], 1u) - tile_base;
  __syncthreads();
  wave_tile<DOWN, true, VIRT>(a, sh.tile, sh, status, state, ntiles, epoch, tot);
}

// The tail of a bucket: once a wave fits one tile every later wave of the bucket does too (wave sizes never
// grow inside a bucket), so ONE block runs them back to back on the device — no host round trip per wave.
// Homopolymer runs produce one tiny wave per base: this is what keeps them cheap.
// out[0] = next wave start p, out[1..4] = cursors, out[5] = waves run.
#define WAVE_TAIL_MAX (WAVE_TILE - 8)
template <bool DOWN>
__global__ void __launch_bounds__(WAVE_THREADS) k_wave_tail(WaveArgs a0, u32 bucket, u64 p0, u32 max_waves,
                                                           u64* __restrict__ out) {
  extern __shared__ __align__(16) unsigned char wave_smem[];
  WaveShared& sh = *reinterpret_cast<WaveShared*>(wave_smem);
  __shared__ u32 s_tot[4];
  if (threadIdx.x == 0) {
    for (int s = 0; s < 4; s++) sh.cur[s] = a0.cur[s];
    sh.p = p0;
    sh.nwaves = 0;
    sh.ntile = 0;
    sh.nmid = 0;
  }
  for (;;) {
    __syncthreads();
    const u64 cc = sh.cur[bucket], p = sh.p;
    const u64 len = DOWN ? p - cc : cc - p;
    if (len == 0 || len > WAVE_TAIL_MAX || sh.nwaves >= max_waves) break;
    if (len > 32 && len <= WAVE_THREADS) {
      // Waves of at most one entry per thread (repeat-rich text: thousands per bucket, shrinking from a few
      // hundred entries to a handful): ballot ranks, one packed count per warp, two block barriers per wave,
      // cursors replicated in registers — and, as in the warp path below, the entries appended to this bucket are
      // handed to the next wave through shared memory instead of a global-memory round trip.  (The full tile
      // machinery — 16 slots per thread, staging, coalesced copy-out — costs ~2 us per wave whatever its size.)
      __syncthreads();  // (every thread has read sh.cur / sh.p)
      const u32 t = threadIdx.x, lane = t & 31u, w = t >> 5;
      u32* nbv = sh.sm.sSA;   // next wave's entries, in rank order
      u32* nbc = sh.sm.sCTX;
      u32* nbr = sh.sm.sSA + 2048;   // ... and what is left of their run bounds
      u32* sv = sh.sm.sSA + 1024;    // the current wave's suffixes by entry index (bursts)
      u64 c0 = sh.cur[0], c1 = sh.cur[1], c2 = sh.cur[2], c3 = sh.cur[3], pp = p;
      u32 nw = sh.nwaves;
      bool have = false;
      for (;;) {
        const u64 cb = bucket == 0 ? c0 : bucket == 1 ? c1 : bucket == 2 ? c2 : c3;
        const u64 ln = DOWN ? pp - cb : cb - pp;
        if (ln <= 32 || ln > WAVE_THREADS || nw >= max_waves) break;  // (uniform: every thread holds the same cursors)
        u32 v = 0, c = 1u, rr = 0;
        if (t < ln) {
          if (have) {
            v = nbv[t];
            c = nbc[t];
            rr = nbr[t];
          } else {
            const u64 slot = DOWN ? pp - 1 - t : pp + t;
            v = a0.SA[slot];
            c = a0.CTX[slot];
          }
        }
        if (c == 1u && v != 0 && v != KISS_EMPTY) c = ctx_gather(a0.T, v);
        const u32 cu = ctx_symbol(c, a0.symmask);
        if (t < ln) rr = (cu == bucket && v != KISS_EMPTY) ? run_bound(a0.T, v, c, rr, bucket) : 0u;
        const u32 wm = __reduce_min_sync(0xffffffffu, t < ln ? rr : 0xFFFFFFFFu);
        const u32 m0 = __ballot_sync(0xffffffffu, cu == 0), m1 = __ballot_sync(0xffffffffu, cu == 1),
                  m2 = __ballot_sync(0xffffffffu, cu == 2), m3 = __ballot_sync(0xffffffffu, cu == 3);
        if (lane == 0) {
          sh.wsum[w] = (u64)__popc(m0) | ((u64)__popc(m1) << 16) | ((u64)__popc(m2) << 32) | ((u64)__popc(m3) << 48);
          sh.wmin[w] = wm;
        }
        sv[t] = v;
        __syncthreads();  // counts published; every thread has read its entry from the hand-over buffer
        u64 before = 0, total = 0;
        u32 K = 0xFFFFFFFFu;
#pragma unroll
        for (u32 ww = 0; ww < WAVE_WARPS; ww++) {
          const u64 x = sh.wsum[ww];
          total += x;
          if (ww < w) before += x;
          K = min(K, sh.wmin[ww]);
        }
        if (K >= BURST_MIN) {
          // every entry has >= K bases of the bucket's symbol before it: waves 1..K are the suffixes v_j - k in the
          // order j of this wave (consecutive threads take consecutive k of one j: adjacent text positions)
          const u32 n32 = (u32)ln, cnt = K * n32;
          for (u32 tt = t; tt < cnt; tt += WAVE_THREADS) {
            const u32 j = tt / K, k = tt - j * K + 1u;
            const u32 val = sv[j] - k;
            const u64 idx = (u64)(k - 1u) * n32 + j;
            const u64 dst = DOWN ? cb - 1 - idx : cb + idx;
            a0.SA[dst] = val;
            a0.CTX[dst] = ctx_gather(a0.T, val);
          }
          if (t < ln) {
            v -= K;
            nbv[t] = v;
            nbc[t] = ctx_gather(a0.T, v);
            nbr[t] = rr - K;
          }
          const u64 adv = (u64)K * n32, last = (u64)(K - 1u) * n32;
          pp = DOWN ? cb - last : cb + last;
          if (DOWN) { c0 -= bucket == 0 ? adv : 0; c1 -= bucket == 1 ? adv : 0; c2 -= bucket == 2 ? adv : 0; c3 -= bucket == 3 ? adv : 0; }
          else { c0 += bucket == 0 ? adv : 0; c1 += bucket == 1 ? adv : 0; c2 += bucket == 2 ? adv : 0; c3 += bucket == 3 ? adv : 0; }
          nw += K;
          have = true;
          __syncthreads();
          continue;
        }
        if (cu < 4) {
          const u32 mm = cu == 0 ? m0 : cu == 1 ? m1 : cu == 2 ? m2 : m3;
          const u64 cs = cu == 0 ? c0 : cu == 1 ? c1 : cu == 2 ? c2 : c3;
          const u32 r = ((u32)(before >> (16 * cu)) & 0xFFFFu) + __popc(mm & ((1u << lane) - 1u));
          const u64 dst = DOWN ? cs - 1 - r : cs + r;
          a0.SA[dst] = v - 1u;
          a0.CTX[dst] = c >> 2;
          if (cu == bucket) { nbv[r] = v - 1u; nbc[r] = c >> 2; nbr[r] = rr ? rr - 1u : 0u; }
        }
        const u32 t0 = (u32)total & 0xFFFFu, t1 = (u32)(total >> 16) & 0xFFFFu, t2 = (u32)(total >> 32) & 0xFFFFu,
                  t3 = (u32)(total >> 48);
        pp = cb;
        if (DOWN) { c0 -= t0; c1 -= t1; c2 -= t2; c3 -= t3; }
        else { c0 += t0; c1 += t1; c2 += t2; c3 += t3; }
        nw++;
        have = true;
        __syncthreads();  // hand-over buffer complete; the counts may be rewritten
      }
      if (t == 0) {
        sh.cur[0] = c0; sh.cur[1] = c1; sh.cur[2] = c2; sh.cur[3] = c3;
        sh.p = pp;
        sh.nmid += nw - sh.nwaves;
        sh.nwaves = nw;
      }
      continue;
    }
    if (len <= 32) {
      // Waves of at most one warp's width (the chains of homopolymer / tandem runs: one wave per base, thousands
      // per bucket on repeat-rich text): warp 0 runs them alone — one entry per lane, ballot ranks, no block
      // barrier, cursors in registers.  A wave never grows inside a bucket, so this normally ends the tail.
      __syncthreads();  // (every thread has read sh.cur / sh.p before warp 0 rewrites them)
      if (threadIdx.x < 32) {
        const u32 lane = threadIdx.x;
        u64 c0 = sh.cur[0], c1 = sh.cur[1], c2 = sh.cur[2], c3 = sh.cur[3], pp = p;
        u32 nw = sh.nwaves;
        // The entries a wave appends to its own bucket ARE the next wave, in rank order: they are handed on in
        // registers (lane r of the next wave takes them from the lane that ranked r), so the chain of waves
        // never waits on a global-memory round trip; only the first wave is loaded.
        bool have = false;
        u32 v = 0, c = 1u, rr = 0;
        for (;;) {
          const u64 cb = bucket == 0 ? c0 : bucket == 1 ? c1 : bucket == 2 ? c2 : c3;
          const u64 ln = DOWN ? pp - cb : cb - pp;
          if (ln == 0 || ln > 32 || nw >= max_waves) break;
          if (!have) {
            v = 0; c = 1u; rr = 0;
            if (lane < ln) {
              const u64 slot = DOWN ? pp - 1 - lane : pp + lane;
              v = a0.SA[slot];
              c = a0.CTX[slot];
            }
          }
          if (c == 1u && v != 0 && v != KISS_EMPTY) c = ctx_gather(a0.T, v);
          const u32 cu = ctx_symbol(c, a0.symmask);
          rr = (lane < ln && cu == bucket && v != KISS_EMPTY) ? run_bound(a0.T, v, c, rr, bucket) : 0u;
          const u32 K = __reduce_min_sync(0xffffffffu, lane < ln ? rr : 0xFFFFFFFFu);
          if (K >= BURST_MIN) {
            // (see the one-entry-per-thread path) waves 1..K in one burst; lanes take consecutive k of one entry j
            const u32 n32 = (u32)ln, cnt = K * n32;
            for (u32 t0 = 0; t0 < cnt; t0 += 32) {
              const u32 tt = t0 + lane, j = tt / K, k = tt - j * K + 1u;
              const u32 vj = __shfl_sync(0xffffffffu, v, j & 31u);
              if (tt < cnt) {
                const u32 val = vj - k;
                const u64 idx = (u64)(k - 1u) * n32 + j;
                const u64 dst = DOWN ? cb - 1 - idx : cb + idx;
                a0.SA[dst] = val;
                a0.CTX[dst] = ctx_gather(a0.T, val);
              }
            }
            if (lane < ln) {
              v -= K;
              rr -= K;
              c = ctx_gather(a0.T, v);
            }
            const u64 adv = (u64)K * n32, last = (u64)(K - 1u) * n32;
            pp = DOWN ? cb - last : cb + last;
            if (DOWN) { c0 -= bucket == 0 ? adv : 0; c1 -= bucket == 1 ? adv : 0; c2 -= bucket == 2 ? adv : 0; c3 -= bucket == 3 ? adv : 0; }
            else { c0 += bucket == 0 ? adv : 0; c1 += bucket == 1 ? adv : 0; c2 += bucket == 2 ? adv : 0; c3 += bucket == 3 ? adv : 0; }
            nw += K;
            have = true;
            continue;
          }
          const u32 m0 = __ballot_sync(0xffffffffu, cu == 0), m1 = __ballot_sync(0xffffffffu, cu == 1),
                    m2 = __ballot_sync(0xffffffffu, cu == 2), m3 = __ballot_sync(0xffffffffu, cu == 3);
          const u32 nv = v - 1u, nc = c >> 2;
          if (cu < 4) {
            const u32 mm = cu == 0 ? m0 : cu == 1 ? m1 : cu == 2 ? m2 : m3;
            const u64 cs = cu == 0 ? c0 : cu == 1 ? c1 : cu == 2 ? c2 : c3;
            const u32 r = __popc(mm & ((1u << lane) - 1u));
            const u64 dst = DOWN ? cs - 1 - r : cs + r;
            a0.SA[dst] = nv;
            a0.CTX[dst] = nc;
          }
          // next wave: lane r <- the lane whose entry ranked r in this bucket
          const u32 mb = bucket == 0 ? m0 : bucket == 1 ? m1 : bucket == 2 ? m2 : m3;
          const u32 src = __fns(mb, 0, lane + 1);  // lane of the (lane+1)-th set bit, 0xffffffff if none
          const u32 tv = __shfl_sync(0xffffffffu, nv, src & 31u), tc = __shfl_sync(0xffffffffu, nc, src & 31u);
          const u32 tr = __shfl_sync(0xffffffffu, rr ? rr - 1u : 0u, src & 31u);
          v = src < 32u ? tv : 0u;
          c = src < 32u ? tc : 1u;
          rr = src < 32u ? tr : 0u;
          have = true;
          pp = cb;
          if (DOWN) { c0 -= __popc(m0); c1 -= __popc(m1); c2 -= __popc(m2); c3 -= __popc(m3); }
          else { c0 += __popc(m0); c1 += __popc(m1); c2 += __popc(m2); c3 += __popc(m3); }
          nw++;
        }
        __syncwarp();
        if (lane == 0) {
          sh.cur[0] = c0; sh.cur[1] = c1; sh.cur[2] = c2; sh.cur[3] = c3;
          sh.p = pp;
          sh.nwaves = nw;
        }
      }
      continue;
    }
    WaveArgs a = a0;
    a.lo = DOWN ? cc : p;
    a.len = len;
    a.cur[0] = sh.cur[0]; a.cur[1] = sh.cur[1]; a.cur[2] = sh.cur[2]; a.cur[3] = sh.cur[3];
    __syncthreads();
    wave_tile<DOWN, false>(a, 0, sh, nullptr, nullptr, 1, 0, s_tot);
    __syncthreads();
    if (threadIdx.x == 0) {
      sh.p = cc;
      for (int s = 0; s < 4; s++) sh.cur[s] = DOWN ? sh.cur[s] - s_tot[s] : sh.cur[s] + s_tot[s];
      sh.nwaves++;
      sh.ntile++;
    }
  }
  if (threadIdx.x == 0) {
    out[0] = sh.p;
    for (int s = 0; s < 4; s++) out[1 + s] = sh.cur[s];
    out[5] = sh.nwaves;
    out[6] = sh.ntile;
    out[7] = sh.nmid;
  }
}

// host side of the tail: returns the number of waves run; p and cur[] are updated
u32 induce_tail(u32* SA, u32* CTX, const u64* T, u32 bucket, u64* p, bool down, u32 symmask, u64 cur[4],
                WaveScratch ws, cudaStream_t st) {
  WaveArgs a;
  a.SA = SA; a.CTX = CTX; a.inSA = SA; a.inCTX = CTX; a.T = T; a.lo = 0; a.len = 0; a.down = down ? 1u : 0u;
  a.shorts = nullptr; a.ins = nullptr; a.ns = 0;
  a.symmask = symmask;
  for (int s = 0; s < 4; s++) a.cur[s] = cur[s];
  {
    ProfScope ps(PK_WAVE_SMALL, st);
    if (down) k_wave_tail<true><<<1, WAVE_THREADS, sizeof(WaveShared), st>>>(a, bucket, *p, 1u << 20, ws.tail_host);
    else k_wave_tail<false><<<1, WAVE_THREADS, sizeof(WaveShared), st>>>(a, bucket, *p, 1u << 20, ws.tail_host);
    CK_LAUNCH(); g_launches++;
  }
  CK(cudaStreamSynchronize(st));
  *p = ws.tail_host[0];
  for (int s = 0; s < 4; s++) cur[s] = ws.tail_host[1 + s];
  static const bool dbg = getenv("KISS_B200_DEBUG_WAVES") != nullptr;
  if (dbg)
    fprintf(stderr, "[kiss_b200] tail bucket %u %s: %llu waves (tile path %llu, one entry per thread %llu, one warp %llu)\n",
            bucket, down ? "S" : "L", (unsigned long long)ws.tail_host[5], (unsigned long long)ws.tail_host[6],
            (unsigned long long)ws.tail_host[7],
            (unsigned long long)(ws.tail_host[5] - ws.tail_host[6] - ws.tail_host[7]));
  return (u32)ws.tail_host[5];
}

void wave_init() {
  const int bytes = (int)sizeof(WaveShared);
  CK(cudaFuncSetAttribute(k_wave_fused<true, WAVE_MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  CK(cudaFuncSetAttribute(k_wave_fused<false, WAVE_MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  CK(cudaFuncSetAttribute(k_wave_fused<false, WAVE_MINB, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  CK(cudaFuncSetAttribute(k_wave_tail<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  CK(cudaFuncSetAttribute(k_wave_tail<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
}

// contexts of a list of suffixes (what the LMS waves would gather themselves): at N > 1 every rank fetches them for
// the key range it sorted, so the 128-byte DRAM fetch per LMS suffix is spread over the GPUs and rank 0's LMS
// waves stream (position, context) pairs
__global__ void __launch_bounds__(256) k_lms_contexts(const u64* __restrict__ T, const u32* __restrict__ pos, u64 cnt,
                                                     u32* __restrict__ out) {
  const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < cnt) out[i] = ctx_gather(T, pos[i]);
}
void lms_contexts(const u64* T, const u32* pos, u64 cnt, u32* out, cudaStream_t st) {
  if (!cnt) return;
  k_lms_contexts<<<(unsigned)((cnt + 255) / 256), 256, 0, st>>>(T, pos, cnt, out);
  CK_LAUNCH(); g_launches++;
}
void induce_wave(u32* SA, u32* CTX, const u64* T, const u32* lms_src, const u32* lms_ctx, u64 lo, u64 len, bool down,
                 u32 symmask, const u64 cur[4], WaveScratch ws, u32 tot[4], cudaStream_t st, const u32* shorts,
                 const u32* ins, u32 ns) {
  tot[0] = tot[1] = tot[2] = tot[3] = 0;
  if (len == 0) return;
  WaveArgs a;
  a.SA = SA; a.CTX = CTX; a.T = T; a.lo = lo; a.len = len; a.down = down ? 1u : 0u; a.symmask = symmask;
  a.inSA = lms_src ? lms_src : SA;
  // (the 16-byte loads of the two scanned arrays share their cut points: a context list must be aligned like its list)
  const bool virt = lms_src && ns && !down;  // the list is (long suffixes, short ones inserted on the fly)
  a.inCTX = lms_src ? ((lms_ctx && (virt || (((uintptr_t)lms_ctx ^ (uintptr_t)lms_src) & 15) == 0)) ? lms_ctx : nullptr) : CTX;
  a.shorts = virt ? shorts : nullptr;
  a.ins = virt ? ins : nullptr;
  a.ns = virt ? ns : 0u;
  for (int s = 0; s < 4; s++) a.cur[s] = cur[s];
  const u64 mis = virt ? 0ull : ((u64)(uintptr_t)a.inSA >> 2) & 3ull, hi = lo + len;
  u64 len_al = down ? (hi + ((4ull - ((hi + mis) & 3ull)) & 3ull) - lo) : (hi - (lo - ((lo + mis) & 3ull)));
  u32 ntiles = (u32)wave_tiles(len_al);
  {
    ProfScope ps(len <= WAVE_TILE ? PK_WAVE_SMALL : (lms_src ? PK_WAVE_LMS : PK_WAVE_SCATTER), st);
    // (a persistent variant that requests the next ticket early measured 2x slower: a ticket taken long before
    // its tile starts makes every later tile wait in the look-back)
    const u32 ep = ++*ws.epoch;
    // (5 resident blocks per SM, 48 registers with a small spill, measured 1 ms slower than 4)
    if (virt) k_wave_fused<false, WAVE_MINB, true><<<ntiles, WAVE_THREADS, sizeof(WaveShared), st>>>(a, ws.status, ws.state, *ws.tile_base, ntiles, ep, ws.tot_host);
    else if (down) k_wave_fused<true, WAVE_MINB><<<ntiles, WAVE_THREADS, sizeof(WaveShared), st>>>(a, ws.status, ws.state, *ws.tile_base, ntiles, ep, ws.tot_host);
    else k_wave_fused<false, WAVE_MINB><<<ntiles, WAVE_THREADS, sizeof(WaveShared), st>>>(a, ws.status, ws.state, *ws.tile_base, ntiles, ep, ws.tot_host);
    CK_LAUNCH(); g_launches++;
    *ws.tile_base += ntiles;
  }
  // (the kernel stored the totals in pinned host memory: no copy engine between the waves and the host)
  CK(cudaStreamSynchronize(st));
  for (int s = 0; s < 4; s++) tot[s] = ws.tot_host[s];
}

__global__ void k_set_sentinel(u32* SA, u32* CTX, u32 n) { SA[0] = n; CTX[0] = 1u; }
void set_sentinel(u32* SA, u32* CTX, u64 n, cudaStream_t st) {
  k_set_sentinel<<<1, 1, 0, st>>>(SA, CTX, (u32)n);
  CK_LAUNCH(); g_launches++;
}

}  // namespace kiss
