// ingest.cu — the step before the path: file bytes -> base codes on the device.
//
// Replaces read_sequence (include/utils/io.hpp:6-18) + the FASTA extraction operator it uses
// (include/biovoltron/file_io/fasta.hpp:116-151) + Codec::to_int and `% 4` (istring.hpp:27-46,
// include/command/suffix_sort.hpp:33).  Semantics kept exactly:
//   * FASTA mode iff the first byte of the file is '>' (io.hpp:8); otherwise text mode: every line is sequence;
//   * lines are what std::getline returns: split at '\n', the '\n' itself dropped, a '\r' before it KEPT
//     (it then encodes like any non-ACGT byte: 4 % 4 = 0 = 'A');
//   * in FASTA mode a line whose FIRST byte is '>' is a header and is dropped whole (fasta.hpp:127, :137-138);
//     a '>' anywhere else is sequence; all records are concatenated without separators (io.hpp:10-11);
//   * a/A c/C g/G t/T -> 0..3, every other byte -> 0.
// The reference does this with serial getline + per-character push_back; here a byte is kept iff it is not '\n' and
// the line it belongs to is not a header, which is a stream compaction:
//   k_ingest_summary  per 4096-byte chunk: kept bytes of the lines that START in the chunk, non-newline bytes of
//                     the leading partial line (kept or not depends on where that line started), last line start
//   k_ingest_bmax / bscan / carry   running maximum of the line starts over the chunks (two levels) -> is the leading
//                     partial line of chunk c a header? -> kept bytes per chunk
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

// exclusive running maximum over the threads of the block (0 = none before)
__device__ __forceinline__ u32 block_excl_max(u32 v, u32* ws) {
  const u32 lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
  u32 incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const u32 x = __shfl_up_sync(0xffffffffu, incl, o);
    if ((int)lane >= o) incl = max(incl, x);
  }
  if (lane == 31) ws[w] = incl;
  __syncthreads();
  u32 before = 0;
  for (u32 ww = 0; ww < w; ww++) before = max(before, ws[ww]);
  u32 ex = __shfl_up_sync(0xffffffffu, incl, 1);
  if (lane == 0) ex = 0;
  __syncthreads();
  return max(before, ex);
}

// Masks of the thread's bytes that belong to a header line (drop) and to the line that started before this chunk
// (lead: kept or not is decided by k_ingest_carry).  Bit-parallel: a header region runs from a header start ('>' at a
// line start) up to the next line start; adding the header-start bits to the word that has a 1 at every header start
// and every byte that does not start a line lets the carry run through exactly those bytes.
__device__ __forceinline__ void ingest_classes(const u8* __restrict__ raw, u64 chunk0, const IngestThread& it, u32 ls_before,
                                               bool fasta, u32& drop, u32& lead) {
  const u32 L = it.lsmask, Hs = fasta ? (L & it.gtmask) : 0u;
  const u32 A = ((~L) | Hs) & 0xFFFFu;
  const u32 S = A + Hs;
  drop = (((S ^ A) | Hs) & A) & it.vmask;
  // the bytes below the first line start of this thread inherit the class of the line they continue
  const u32 P = (L ? ((L & (0u - L)) - 1u) : 0xFFFFu) & it.vmask;
  lead = 0;
  if (P) {
    if (!ls_before) lead = P;
    else if (fasta && raw[chunk0 + ls_before - 1] == '>') drop |= P;
  }
}

__global__ void __launch_bounds__(ING_THREADS) k_ingest_summary(const u8* __restrict__ raw, u64 nbytes, bool fasta,
                                                               u32* __restrict__ lead, u32* __restrict__ rest,
                                                               u64* __restrict__ last_ls) {
  __shared__ u32 ws[ING_THREADS / 32];
  __shared__ u32 s_lead, s_rest, s_last;
  if (threadIdx.x == 0) { s_lead = 0; s_rest = 0; s_last = 0; }
  const u64 chunk0 = (u64)blockIdx.x * ING_CHUNK;
  IngestThread it;
  ingest_load(raw, nbytes, chunk0, it);
  const u32 ls_before = block_excl_max(it.last_ls, ws);  // (its barriers also order the zeroing above)
  u32 drop, leadm;
  ingest_classes(raw, chunk0, it, ls_before, fasta, drop, leadm);
  const u32 body = it.vmask & ~it.nlmask;
  u32 nlead = (u32)__popc(body & leadm), nkeep = (u32)__popc(body & ~leadm & ~drop);
  nlead = __reduce_add_sync(0xffffffffu, nlead);
  nkeep = __reduce_add_sync(0xffffffffu, nkeep);
  const u32 lmax = __reduce_max_sync(0xffffffffu, it.last_ls);
  if ((threadIdx.x & 31u) == 0) {
    if (nlead) atomicAdd(&s_lead, nlead);
    if (nkeep) atomicAdd(&s_rest, nkeep);
    if (lmax) atomicMax(&s_last, lmax);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    lead[blockIdx.x] = s_lead;
    rest[blockIdx.x] = s_rest;
    last_ls[blockIdx.x] = s_last ? chunk0 + s_last : 0ull;  // file index + 1
  }
}

// Header flag of every chunk's leading partial line and kept bytes per chunk (into `rest`): the line containing a
// chunk's first byte starts at the running maximum of the line starts of the chunks before it.  Two levels: blocks of
// 1024 chunks publish their maximum (k_ingest_bmax), one block scans those (k_ingest_bscan), every block then scans
// its own 1024 chunks (k_ingest_carry).  (One block walking all chunks took 0.9 ms per GB of file.)
__device__ __forceinline__ u64 block_excl_max64(u64 v, u64* ws, u64* total) {
  const u32 lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
  u64 incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const u64 x = __shfl_up_sync(0xffffffffu, incl, o);
    if ((int)lane >= o) incl = max(incl, x);
  }
  if (lane == 31) ws[w] = incl;
  __syncthreads();
  u64 before = 0, all = 0;
  for (u32 ww = 0; ww < 32; ww++) {
    const u64 x = ws[ww];
    if (ww < w) before = max(before, x);
    all = max(all, x);
  }
  u64 ex = __shfl_up_sync(0xffffffffu, incl, 1);
  if (lane == 0) ex = 0;
  __syncthreads();
  if (total) *total = all;
  return max(before, ex);
}

__global__ void __launch_bounds__(1024) k_ingest_bmax(const u64* __restrict__ last_ls, u64 nchunks, u64* __restrict__ bmax) {
  __shared__ u64 ws[32];
  const u64 c = (u64)blockIdx.x * 1024 + threadIdx.x;
  u64 all;
  block_excl_max64(c < nchunks ? last_ls[c] : 0ull, ws, &all);
  if (threadIdx.x == 0) bmax[blockIdx.x] = all;
}

// in place: bmax[b] <- maximum over the blocks before b; *total zeroed for the kept-byte sum
__global__ void __launch_bounds__(1024) k_ingest_bscan(u64* __restrict__ bmax, u64 nblocks, u64* __restrict__ total) {
  __shared__ u64 ws[32];
  __shared__ u64 s_carry;
  if (threadIdx.x == 0) { s_carry = 0; *total = 0; }
  __syncthreads();
  for (u64 base = 0; base < nblocks; base += 1024) {
    const u64 i = base + threadIdx.x;
    u64 all;
    const u64 ex = block_excl_max64(i < nblocks ? bmax[i] : 0ull, ws, &all);
    const u64 carry = s_carry;
    if (i < nblocks) bmax[i] = max(carry, ex);
    __syncthreads();
    if (threadIdx.x == 0) s_carry = max(carry, all);
    __syncthreads();
  }
}

__global__ void __launch_bounds__(1024) k_ingest_carry(const u8* __restrict__ raw, bool fasta, u64 nchunks,
                                                      const u32* __restrict__ lead, u32* __restrict__ rest,
                                                      const u64* __restrict__ last_ls, const u64* __restrict__ bmax,
                                                      u8* __restrict__ hflag, u64* __restrict__ total) {
  __shared__ u64 ws[32];
  __shared__ unsigned long long s_total;
  if (threadIdx.x == 0) s_total = 0ull;
  const u64 c = (u64)blockIdx.x * 1024 + threadIdx.x;
  const u64 run = max(bmax[blockIdx.x], block_excl_max64(c < nchunks ? last_ls[c] : 0ull, ws, nullptr));
  u32 kept = 0;
  if (c < nchunks) {
    const bool h = fasta && run != 0 && raw[run - 1] == '>';
    kept = rest[c] + (h ? 0u : lead[c]);
    hflag[c] = h ? 1 : 0;
    rest[c] = kept;
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
  const u32 ls_before = block_excl_max(it.last_ls, ws);
  u32 drop, leadm;
  ingest_classes(raw, chunk0, it, ls_before, fasta, drop, leadm);
  if (hflag[blockIdx.x]) drop |= leadm;
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

// codes_dev needs room for nbytes bytes.  lead / rest: nchunks u32, last_ls: nchunks u64, bmax: nchunks / 1024 + 1 u64,
// hflag: nchunks bytes,
// partials: scan_partials_len(nchunks), total_dev: one u64.  Everything is enqueued on st; *total_dev holds n afterwards.
void ingest(const u8* raw_dev, u64 nbytes, u8* codes_dev, u32* lead, u32* rest, u64* last_ls, u64* bmax, u8* hflag,
            u32* partials, u64* total_dev, bool fasta, cudaStream_t st) {
  const u64 nch = ingest_chunks(nbytes);
  if (!nch) return;
  k_ingest_summary<<<(unsigned)nch, ING_THREADS, 0, st>>>(raw_dev, nbytes, fasta, lead, rest, last_ls);
  CK_LAUNCH();
  const u64 nb = (nch + 1023) / 1024;
  k_ingest_bmax<<<(unsigned)nb, 1024, 0, st>>>(last_ls, nch, bmax);
  CK_LAUNCH();
  k_ingest_bscan<<<1, 1024, 0, st>>>(bmax, nb, total_dev);
  CK_LAUNCH();
  k_ingest_carry<<<(unsigned)nb, 1024, 0, st>>>(raw_dev, fasta, nch, lead, rest, last_ls, bmax, hflag, total_dev);
  CK_LAUNCH();
  scan_excl_u32(rest, nch, partials, nullptr, st);
  k_ingest_emit<<<(unsigned)nch, ING_THREADS, 0, st>>>(raw_dev, nbytes, fasta, rest, hflag, 0ull, codes_dev);
  CK_LAUNCH();
  g_launches += 5;
}

}  // namespace kiss
