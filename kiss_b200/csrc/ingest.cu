// ingest.cu — the step before the path: file bytes -> base codes on the device.
//
// Replaces read_sequence (include/utils/io.hpp:6-18) + the FASTA extraction operator it uses
// (include/biovoltron/file_io/fasta.hpp:116-151) + Codec::to_int and `% 4` (istring.hpp:27-46,
// include/command/suffix_sort.hpp:33).  Semantics kept exactly:
//   * FASTA mode iff the first byte of the file is '>' (io.hpp:8); otherwise text mode: every line is sequence;
//   * lines are what std::getline returns: split at '\n', the '\n' itself dropped, a '\r' before it KEPT
//     (it then encodes like any non-ACGT byte: 4 % 4 = 0 = 'A');
//   * in FASTA mode a line whose FIRST byte is '>' is a header and is dropped whole (fasta.hpp:127, :137-138) — unless
//     the line before it was a header: the line after a header is sequence whatever it starts with (:129-133), so
//     header-looking lines in a row alternate; a '>' anywhere else is sequence; all records are concatenated without
//     separators (io.hpp:10-11);
//   * a/A c/C g/G t/T -> 0..3, every other byte -> 0.
// The reference does this with serial getline + per-character push_back; here a byte is kept iff it is not '\n' and
// the line it belongs to is not a header, which is a stream compaction:
//   k_ingest_summary  per 4096-byte chunk: what the chunk does to the one-bit state "the current line is a header"
//                     (one of four functions) and the bytes it keeps under either incoming state
//   k_ingest_bfn / bscan / carry   scan of those functions over the chunks (two levels) -> state at every chunk's
//                     first byte -> kept bytes per chunk
//   (exclusive scan of the kept counts, scan.cu)
//   k_ingest_emit     per chunk: flags again, block scan, codes written at the chunk's offset
// Algorithmic bytes: 2 reads of the file + n written.
#include "common.cuh"
#include "kernels.h"

namespace kiss {

#define ING_THREADS 256
#define ING_PER 16
#define ING_CHUNK (ING_THREADS * ING_PER)

__device__ __forceinline__ u32 ingest_code(u32 c) {
  c |= 0x20u;  // lower case
  return c == 'c' ? 1u : c == 'g' ? 2u : c == 't' ? 3u : 0u;
}

// classes of a byte's line
#define ING_LEAD 0u  // the line started before this chunk
#define ING_KEEP 1u
#define ING_DROP 2u  // header line

struct IngestThread {
  u32 w[4];      // the thread's 16 bytes, byte k in bits 8 (k & 3) .. of word k / 4
  u32 valid;     // number of bytes of this thread inside the file
  u32 vmask;     // bit k = byte k is inside the file
  u32 nlmask;    // byte == '\n'
  u32 gtmask;    // byte == '>'
  u32 lsmask;    // byte is the first of a line
  u32 last_ls;   // chunk-local index + 1 of the last line start among this thread's bytes, 0 = none
};

// bit k of the result = byte k of w equals c (c replicated in all four bytes of c4)
__device__ __forceinline__ u32 eq_mask4(u32 w, u32 c4) {
  return ((__vcmpeq4(w, c4) & 0x01010101u) * 0x01020408u) >> 24;  // bytes 0..3 -> bits 0..3
}

__device__ __forceinline__ void ingest_load(const u8* __restrict__ raw, u64 nbytes, u64 chunk0, IngestThread& it) {
  const u64 i0 = chunk0 + (u64)threadIdx.x * ING_PER;
  it.valid = i0 < nbytes ? (u32)min((u64)ING_PER, nbytes - i0) : 0u;
  it.vmask = it.valid >= 16 ? 0xFFFFu : ((1u << it.valid) - 1u);
  if (it.valid == ING_PER && (((uintptr_t)(raw + i0)) & 15) == 0) {
    const uint4 q = __ldg(reinterpret_cast<const uint4*>(raw + i0));
    it.w[0] = q.x; it.w[1] = q.y; it.w[2] = q.z; it.w[3] = q.w;
  } else {
#pragma unroll
    for (int q = 0; q < 4; q++) {
      u32 x = 0;
#pragma unroll
      for (int k = 0; k < 4; k++)
        if ((u32)(4 * q + k) < it.valid) x |= (u32)raw[i0 + 4 * q + k] << (8 * k);
      it.w[q] = x;
    }
  }
  const bool prev_nl = i0 == 0 || (it.valid && raw[i0 - 1] == '\n');
  u32 nl = 0, gt = 0;
#pragma unroll
  for (int q = 0; q < 4; q++) {
    nl |= eq_mask4(it.w[q], 0x0A0A0A0Au) << (4 * q);
    gt |= eq_mask4(it.w[q], 0x3E3E3E3Eu) << (4 * q);
  }
  it.nlmask = nl & it.vmask;
  it.gtmask = gt & it.vmask;
  it.lsmask = ((it.nlmask << 1) | (prev_nl ? 1u : 0u)) & it.vmask;
  it.last_ls = it.lsmask ? threadIdx.x * ING_PER + (32u - (u32)__clz(it.lsmask)) : 0u;
}

// ---- header state.  "Is the current line a header" is a one-bit state that every line start updates:
// new = (the line starts with '>') && !old — the line after a header is sequence whatever it starts with
// (fasta.hpp:129-133), so header-looking lines in a row alternate.  The effect of any stretch of the file on that bit
// is one of the four functions {0,1} -> {0,1}, kept as two bits (bit s = image of s), and stretches compose
// associatively: the state at every thread, chunk and block of chunks comes from scans of such functions.
#define ING_ID 2u  // identity: 0 -> 0, 1 -> 1
__device__ __forceinline__ u32 fn_apply(u32 f, u32 s) { return (f >> s) & 1u; }
__device__ __forceinline__ u32 fn_then(u32 earlier, u32 later) {  // later o earlier
  return fn_apply(later, fn_apply(earlier, 0u)) | (fn_apply(later, fn_apply(earlier, 1u)) << 1);
}
// the thread's 16 bytes under incoming state s: mask of the bytes that belong to header lines, and the state after them
__device__ __forceinline__ u32 ingest_walk(const IngestThread& it, bool fasta, u32 s, u32& drop) {
  u32 L = it.lsmask, prev = 0;
  drop = 0;
  while (L) {
    const u32 k = (u32)__ffs((int)L) - 1u;
    L &= L - 1u;
    if (s) drop |= ((1u << k) - 1u) & ~((1u << prev) - 1u);
    s = (fasta && ((it.gtmask >> k) & 1u) && !s) ? 1u : 0u;
    prev = k;
  }
  if (s) drop |= 0xFFFFu & ~((1u << prev) - 1u);
  drop &= it.vmask;
  return s;
}
__device__ __forceinline__ u32 ingest_fn(const IngestThread& it, bool fasta) {
  if (!it.lsmask) return ING_ID;
  u32 d;
  return ingest_walk(it, fasta, 0u, d) | (ingest_walk(it, fasta, 1u, d) << 1);
}
// exclusive scan of the threads' functions over the block (thread 0 gets the identity); *total = the whole block's
__device__ __forceinline__ u32 block_excl_fn(u32 f, u32* ws, u32* total) {
  const u32 lane = threadIdx.x & 31u, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  u32 incl = f;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const u32 x = __shfl_up_sync(0xffffffffu, incl, o);
    if ((int)lane >= o) incl = fn_then(x, incl);
  }
  if (lane == 31) ws[w] = incl;
  __syncthreads();
  u32 before = ING_ID, all = ING_ID;
  for (u32 ww = 0; ww < nw; ww++) {
    if (ww < w) before = fn_then(before, ws[ww]);
    all = fn_then(all, ws[ww]);
  }
  u32 ex = __shfl_up_sync(0xffffffffu, incl, 1);
  if (lane == 0) ex = ING_ID;
  __syncthreads();
  if (total) *total = all;
  return fn_then(before, ex);
}

// per chunk: its function and the bytes it keeps under either incoming state
__global__ void __launch_bounds__(ING_THREADS) k_ingest_summary(const u8* __restrict__ raw, u64 nbytes, bool fasta,
                                                               u32* __restrict__ kept0, u32* __restrict__ kept1,
                                                               u8* __restrict__ fn) {
  __shared__ u32 ws[ING_THREADS / 32];
  __shared__ u32 s_k0, s_k1;
  if (threadIdx.x == 0) { s_k0 = 0; s_k1 = 0; }
  const u64 chunk0 = (u64)blockIdx.x * ING_CHUNK;
  IngestThread it;
  ingest_load(raw, nbytes, chunk0, it);
  u32 all;
  const u32 ex = block_excl_fn(ingest_fn(it, fasta), ws, &all);  // (its barriers also order the zeroing above)
  const u32 body = it.vmask & ~it.nlmask;
  u32 d0, d1;
  ingest_walk(it, fasta, fn_apply(ex, 0u), d0);
  if (fn_apply(ex, 1u) == fn_apply(ex, 0u)) d1 = d0; else ingest_walk(it, fasta, fn_apply(ex, 1u), d1);
  const u32 n0 = __reduce_add_sync(0xffffffffu, (u32)__popc(body & ~d0));
  const u32 n1 = __reduce_add_sync(0xffffffffu, (u32)__popc(body & ~d1));
  if ((threadIdx.x & 31u) == 0) {
    if (n0) atomicAdd(&s_k0, n0);
    if (n1) atomicAdd(&s_k1, n1);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    kept0[blockIdx.x] = s_k0;
    kept1[blockIdx.x] = s_k1;
    fn[blockIdx.x] = (u8)all;
  }
}

// State at the start of every chunk, two levels: blocks of 1024 chunks publish their function (k_ingest_bfn), one block
// scans those (k_ingest_bscan), every block then scans its own 1024 chunks and picks the kept count of each
// (k_ingest_carry).  (One block walking all chunks took 0.9 ms per GB of file.)
__global__ void __launch_bounds__(1024) k_ingest_bfn(const u8* __restrict__ fn, u64 nchunks, u8* __restrict__ bfn) {
  __shared__ u32 ws[32];
  const u64 c = (u64)blockIdx.x * 1024 + threadIdx.x;
  u32 all;
  block_excl_fn(c < nchunks ? (u32)fn[c] : ING_ID, ws, &all);
  if (threadIdx.x == 0) bfn[blockIdx.x] = (u8)all;
}

// in place: bfn[b] <- state at the start of block b (the file starts outside a header); *total zeroed
__global__ void __launch_bounds__(1024) k_ingest_bscan(u8* __restrict__ bfn, u64 nblocks, u64* __restrict__ total) {
  __shared__ u32 ws[32];
  __shared__ u32 s_carry;
  if (threadIdx.x == 0) { s_carry = ING_ID; *total = 0; }
  __syncthreads();
  for (u64 base = 0; base < nblocks; base += 1024) {
    const u64 i = base + threadIdx.x;
    u32 all;
    const u32 ex = block_excl_fn(i < nblocks ? (u32)bfn[i] : ING_ID, ws, &all);
    const u32 carry = s_carry;
    if (i < nblocks) bfn[i] = (u8)fn_apply(fn_then(carry, ex), 0u);
    __syncthreads();
    if (threadIdx.x == 0) s_carry = fn_then(carry, all);
    __syncthreads();
  }
}

__global__ void __launch_bounds__(1024) k_ingest_carry(u64 nchunks, u32* __restrict__ kept0, const u32* __restrict__ kept1,
                                                      const u8* __restrict__ fn, const u8* __restrict__ bstate,
                                                      u8* __restrict__ hflag, u64* __restrict__ total) {
  __shared__ u32 ws[32];
  __shared__ unsigned long long s_total;
  if (threadIdx.x == 0) s_total = 0ull;
  const u64 c = (u64)blockIdx.x * 1024 + threadIdx.x;
  const u32 ex = block_excl_fn(c < nchunks ? (u32)fn[c] : ING_ID, ws, nullptr);
  u32 kept = 0;
  if (c < nchunks) {
    const u32 s = fn_apply(ex, (u32)bstate[blockIdx.x]);
    kept = s ? kept1[c] : kept0[c];
    hflag[c] = (u8)s;
    kept0[c] = kept;
  }
  kept = __reduce_add_sync(0xffffffffu, kept);
  if ((threadIdx.x & 31u) == 0 && kept) atomicAdd(&s_total, (unsigned long long)kept);
  __syncthreads();
  if (threadIdx.x == 0 && s_total) atomicAdd((unsigned long long*)total, s_total);
}

__global__ void __launch_bounds__(ING_THREADS) k_ingest_emit(const u8* __restrict__ raw, u64 nbytes, bool fasta,
                                                            const u32* __restrict__ offset, const u8* __restrict__ hflag,
                                                            u64 base_off_hi, u8* __restrict__ codes) {
  __shared__ u32 ws[ING_THREADS / 32 + 1];
  __shared__ u32 ws2[33];
  const u64 chunk0 = (u64)blockIdx.x * ING_CHUNK;
  IngestThread it;
  ingest_load(raw, nbytes, chunk0, it);
  const u32 ex = block_excl_fn(ingest_fn(it, fasta), ws, nullptr);
  u32 drop;
  ingest_walk(it, fasta, fn_apply(ex, (u32)hflag[blockIdx.x]), drop);  // hflag = state at the chunk's first byte
  const u32 keep = it.vmask & ~it.nlmask & ~drop;
  // exclusive scan of the per-thread counts
  const u32 cnt = (u32)__popc(keep), lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
  u32 incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const u32 x = __shfl_up_sync(0xffffffffu, incl, o);
    if ((int)lane >= o) incl += x;
  }
  if (lane == 31) ws2[w] = incl;
  __syncthreads();
  u32 before = 0, total = 0;
  for (u32 ww = 0; ww < ING_THREADS / 32; ww++) {
    if (ww < w) before += ws2[ww];
    total += ws2[ww];
  }
  // staged in shared memory, then written with consecutive threads on consecutive bytes (a thread's own 16 codes
  // start at an arbitrary byte: storing them directly costs a partial sector per byte store)
  __shared__ __align__(16) u8 sout[ING_CHUNK];
  u32 lp = before + incl - cnt;
  u32 cw[4];  // codes of the 16 bytes, four per word: a/A c/C g/G t/T -> 0..3, everything else 0
#pragma unroll
  for (int q = 0; q < 4; q++) {
    const u32 lw = it.w[q] | 0x20202020u;
    cw[q] = (__vcmpeq4(lw, 0x63636363u) & 0x01010101u) | (__vcmpeq4(lw, 0x67676767u) & 0x02020202u) |
            (__vcmpeq4(lw, 0x74747474u) & 0x03030303u);
  }
  if (keep == 0xFFFFu && (lp & 3u) == 0) {  // the common case: 16 bases, word-aligned in the staging buffer
    u32* so = reinterpret_cast<u32*>(sout + lp);
    so[0] = cw[0]; so[1] = cw[1]; so[2] = cw[2]; so[3] = cw[3];
  } else {
#pragma unroll
    for (int k = 0; k < ING_PER; k++)
      if ((keep >> k) & 1u) sout[lp++] = (u8)(cw[k >> 2] >> (8 * (k & 3)));
  }
  __syncthreads();
  u8* dst = codes + base_off_hi + offset[blockIdx.x];
  const u32 head = min(total, (u32)((16 - ((uintptr_t)dst & 15)) & 15));  // bytes before the first 16-byte boundary
  if (threadIdx.x < head) dst[threadIdx.x] = sout[threadIdx.x];
  const u32 nvec = (total - head) / 16;
  for (u32 v = threadIdx.x; v < nvec; v += ING_THREADS) {
    const u8* sp = sout + head + 16 * v;
    u32 w4[4];
#pragma unroll
    for (int q = 0; q < 4; q++) w4[q] = (u32)sp[4 * q] | ((u32)sp[4 * q + 1] << 8) | ((u32)sp[4 * q + 2] << 16) | ((u32)sp[4 * q + 3] << 24);
    *reinterpret_cast<uint4*>(dst + head + 16 * v) = make_uint4(w4[0], w4[1], w4[2], w4[3]);
  }
  const u32 tail0 = head + 16 * nvec;
  if (tail0 + threadIdx.x < total) dst[tail0 + threadIdx.x] = sout[tail0 + threadIdx.x];
}

u64 ingest_chunks(u64 nbytes) { return (nbytes + ING_CHUNK - 1) / ING_CHUNK; }

// codes_dev needs room for nbytes bytes.  kept0 / kept1: nchunks u32, fn / hflag: nchunks bytes, bfn: nchunks / 1024 + 1
// bytes, partials: scan_partials_len(nchunks), total_dev: one u64.  Everything is enqueued on st; *total_dev holds n
// afterwards.
void ingest(const u8* raw_dev, u64 nbytes, u8* codes_dev, u32* kept0, u32* kept1, u8* fn, u8* bfn, u8* hflag,
            u32* partials, u64* total_dev, bool fasta, cudaStream_t st) {
  const u64 nch = ingest_chunks(nbytes);
  if (!nch) return;
  u32* rest = kept0;
  k_ingest_summary<<<(unsigned)nch, ING_THREADS, 0, st>>>(raw_dev, nbytes, fasta, kept0, kept1, fn);
  CK_LAUNCH();
  const u64 nb = (nch + 1023) / 1024;
  k_ingest_bfn<<<(unsigned)nb, 1024, 0, st>>>(fn, nch, bfn);
  CK_LAUNCH();
  k_ingest_bscan<<<1, 1024, 0, st>>>(bfn, nb, total_dev);
  CK_LAUNCH();
  k_ingest_carry<<<(unsigned)nb, 1024, 0, st>>>(nch, kept0, kept1, fn, bfn, hflag, total_dev);
  CK_LAUNCH();
  scan_excl_u32(rest, nch, partials, nullptr, st);
  k_ingest_emit<<<(unsigned)nch, ING_THREADS, 0, st>>>(raw_dev, nbytes, fasta, rest, hflag, 0ull, codes_dev);
  CK_LAUNCH();
  g_launches += 5;
}

}  // namespace kiss
