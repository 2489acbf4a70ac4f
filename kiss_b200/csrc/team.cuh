// team.cuh — register-resident bitonic sort of a team (a warp, or the warps of a block) over 64- or 128-bit keys:
// shared by the tie-group teams of lms_sort.cu and the segmented sorts of the doubling rounds (pd.cu).
#pragma once
#include "common.cuh"

#define MK_SLOT_BITS 11

template <int NW>
__device__ __forceinline__ void team_sync() {
  if (NW == 1) __syncwarp(); else __syncthreads();
}
template <int NW>
__device__ __forceinline__ bool team_or(bool v) {
  if (NW == 1) return __any_sync(0xffffffffu, v);
  return __syncthreads_or(v);
}


// order of two keys (never equal: the slot field is unique)
template <bool WIDE>
__device__ __forceinline__ bool mk_less(u64 ah, u64 al, u64 bh, u64 bl) {
  if constexpr (!WIDE) return ah < bh;
  const u64 ka = ah >> MK_SLOT_BITS, kb = bh >> MK_SLOT_BITS;
  return ka != kb ? ka < kb : (al != bl ? al < bl : ah < bh);
}

// ascending bitonic sort of the team's NW * R * 32 keys (slot order = (w R + r) 32 + lane).  The stage (k2) and
// shuffle / shared-memory step (j) loops stay ROLLED: one body per kind of step — a fully unrolled network is 66
// steps x R at 2048 slots, 44 k instructions that miss the instruction cache (12 % of the stall samples) and made
// ptxas keep x[] in local memory (10 % of all instructions were LDL / STL, profiles/r1_c3_medium2048_v0.txt).
// Only the register steps (partner = another register of the same lane) need a compile-time distance.
template <int NW, int R, int JR, bool WIDE>
__device__ __forceinline__ void sort_step_reg(u64 (&xh)[R], u64 (&xl)[R], const u32 w, const u32 lane, const u32 k2) {
#pragma unroll
  for (int r = 0; r < R; r++) {
    if ((r & JR) == 0) {
      const u32 i = ((w * R + r) << 5) | lane;
      const bool asc = (i & k2) == 0;
      const bool swap = mk_less<WIDE>(xh[r], xl[r], xh[r | JR], xl[r | JR]) != asc;
      const u64 ah = xh[r], bh = xh[r | JR];
      xh[r] = swap ? bh : ah;
      xh[r | JR] = swap ? ah : bh;
      if constexpr (WIDE) {
        const u64 al = xl[r], bl = xl[r | JR];
        xl[r] = swap ? bl : al;
        xl[r | JR] = swap ? al : bl;
      }
    }
  }
}
template <int NW, int R, bool WIDE>
__device__ __forceinline__ void team_sort(u64 (&xh)[R], u64 (&xl)[R], const u32 w, const u32 lane, u64* sxh, u64* sxl) {
  constexpr u32 P = NW * R * 32;
  constexpr int UNROLL = P <= 64 ? 16 : 1;  // the small warp teams keep the unrolled network (measured faster there)
#pragma unroll UNROLL
  for (u32 k2 = 2; k2 <= P; k2 <<= 1) {
    if constexpr (NW > 1) {  // partner in another warp of the block
#pragma unroll 1
      for (u32 j = k2 >> 1; j >= 32u * R; j >>= 1) {
#pragma unroll
        for (int r = 0; r < R; r++) {
          sxh[((w * R + r) << 5) | lane] = xh[r];
          if constexpr (WIDE) sxl[((w * R + r) << 5) | lane] = xl[r];
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < R; r++) {
          const u32 i = ((w * R + r) << 5) | lane;
          const u64 yh = sxh[i ^ j];
          u64 yl = 0;
          if constexpr (WIDE) yl = sxl[i ^ j];
          const bool takemin = ((i & k2) == 0) == ((i & j) == 0);
          const bool keep = mk_less<WIDE>(xh[r], xl[r], yh, yl) == takemin;
          xh[r] = keep ? xh[r] : yh;
          if constexpr (WIDE) xl[r] = keep ? xl[r] : yl;
        }
        __syncthreads();
      }
    }
    if constexpr (R > 16) { if (k2 > 32u * 16) sort_step_reg<NW, R, 16, WIDE>(xh, xl, w, lane, k2); }
    if constexpr (R > 8) { if (k2 > 32u * 8) sort_step_reg<NW, R, 8, WIDE>(xh, xl, w, lane, k2); }
    if constexpr (R > 4) { if (k2 > 32u * 4) sort_step_reg<NW, R, 4, WIDE>(xh, xl, w, lane, k2); }
    if constexpr (R > 2) { if (k2 > 32u * 2) sort_step_reg<NW, R, 2, WIDE>(xh, xl, w, lane, k2); }
    if constexpr (R > 1) { if (k2 > 32u) sort_step_reg<NW, R, 1, WIDE>(xh, xl, w, lane, k2); }
#pragma unroll UNROLL
    for (u32 j = k2 > 32u ? 16u : (k2 >> 1); j > 0; j >>= 1) {
#pragma unroll
      for (int r = 0; r < R; r++) {
        const u32 i = ((w * R + r) << 5) | lane;
        const u64 yh = __shfl_xor_sync(0xffffffffu, xh[r], j);
        u64 yl = 0;
        if constexpr (WIDE) yl = __shfl_xor_sync(0xffffffffu, xl[r], j);
        const bool takemin = ((i & k2) == 0) == ((lane & j) == 0);
        const bool keep = mk_less<WIDE>(xh[r], xl[r], yh, yl) == takemin;
        xh[r] = keep ? xh[r] : yh;
        if constexpr (WIDE) xl[r] = keep ? xl[r] : yl;
      }
    }
  }
}
