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
//   k_ingest_carry    one block: running maximum of the line starts over the chunks -> is the leading partial line
//                     of chunk c a header? -> kept bytes per chunk
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
  u8 b[ING_PER];
  u32 valid;     // number of bytes of this thread inside the file
  u32 nlmask;    // byte == '\n'
  u32 lsmask;    // byte is the first of a line
  u32 last_ls;   // chunk-local index + 1 of the last line start among this thread's bytes, 0 = none
};

__device__ __forceinline__ void ingest_load(const u8* __restrict__ raw, u64 nbytes, u64 chunk0, IngestThread& it) {
  const u64 i0 = chunk0 + (u64)threadIdx.x * ING_PER;
  it.valid = i0 < nbytes ? (u32)min((u64)ING_PER, nbytes - i0) : 0u;
  if (it.valid == ING_PER && (((uintptr_t)(raw + i0)) & 15) == 0) {
    const uint4 q = __ldg(reinterpret_cast<const uint4*>(raw + i0));
    const u32 w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int k = 0; k < ING_PER; k++) it.b[k] = (u8)(w[k >> 2] >> (8 * (k & 3)));
  } else {
#pragma unroll
    for (int k = 0; k < ING_PER; k++) it.b[k] = (u32)k < it.valid ? raw[i0 + k] : (u8)0;
  }
  const bool prev_nl = i0 == 0 || (it.valid && raw[i0 - 1] == '\n');
  it.nlmask = 0; it.lsmask = 0; it.last_ls = 0;
#pragma unroll
  for (int k = 0; k < ING_PER; k++) {
    if ((u32)k < it.valid) {
      const bool ls = k == 0 ? prev_nl : it.b[k - 1] == '\n';
      if (it.b[k] == '\n') it.nlmask |= 1u << k;
      if (ls) { it.lsmask |= 1u << k; it.last_ls = threadIdx.x * ING_PER + k + 1; }
    }
  }
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

// class of every byte of the thread: cls bits 2k..2k+1
__device__ __forceinline__ u32 ingest_classes(const u8* __restrict__ raw, u64 chunk0, const IngestThread& it, u32 ls_before,
                                              bool fasta) {
  u32 cur = ING_LEAD;
  if (ls_before) cur = (fasta && raw[chunk0 + ls_before - 1] == '>') ? ING_DROP : ING_KEEP;
  u32 cls = 0;
#pragma unroll
  for (int k = 0; k < ING_PER; k++) {
    if ((it.lsmask >> k) & 1u) cur = (fasta && it.b[k] == '>') ? ING_DROP : ING_KEEP;
    cls |= cur << (2 * k);
  }
  return cls;
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
  const u32 cls = ingest_classes(raw, chunk0, it, ls_before, fasta);
  u32 nlead = 0, nkeep = 0;
#pragma unroll
  for (int k = 0; k < ING_PER; k++) {
    if ((u32)k < it.valid && !((it.nlmask >> k) & 1u)) {
      const u32 c = (cls >> (2 * k)) & 3u;
      nlead += c == ING_LEAD;
      nkeep += c == ING_KEEP;
    }
  }
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

// one block: header flag of every chunk's leading partial line, kept bytes per chunk (into `rest`), grand total
__global__ void __launch_bounds__(1024) k_ingest_carry(const u8* __restrict__ raw, bool fasta, u64 nchunks,
                                                      const u32* __restrict__ lead, u32* __restrict__ rest,
                                                      const u64* __restrict__ last_ls, u8* __restrict__ hflag,
                                                      u64* __restrict__ total) {
  __shared__ u64 smax[1024];
  __shared__ unsigned long long s_total;
  if (threadIdx.x == 0) s_total = 0ull;
  const u64 per = (nchunks + 1023) / 1024;
  const u64 c0 = min(nchunks, (u64)threadIdx.x * per), c1 = min(nchunks, c0 + per);
  u64 m = 0;
  for (u64 c = c0; c < c1; c++) m = max(m, last_ls[c]);
  smax[threadIdx.x] = m;
  __syncthreads();
  if (threadIdx.x == 0) {  // exclusive running maximum over the 1024 ranges
    u64 run = 0;
    for (int t = 0; t < 1024; t++) {
      const u64 x = smax[t];
      smax[t] = run;
      run = max(run, x);
    }
  }
  __syncthreads();
  u64 run = smax[threadIdx.x], sum = 0;
  for (u64 c = c0; c < c1; c++) {
    const bool h = fasta && run != 0 && raw[run - 1] == '>';
    const u32 kept = rest[c] + (h ? 0u : lead[c]);
    hflag[c] = h ? 1 : 0;
    rest[c] = kept;
    sum += kept;
    run = max(run, last_ls[c]);
  }
  if (sum) atomicAdd(&s_total, (unsigned long long)sum);
  __syncthreads();
  if (threadIdx.x == 0) *total = s_total;
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
  const u32 cls = ingest_classes(raw, chunk0, it, ls_before, fasta);
  const u32 lead_cls = hflag[blockIdx.x] ? ING_DROP : ING_KEEP;
  u32 keep = 0;
#pragma unroll
  for (int k = 0; k < ING_PER; k++) {
    if ((u32)k < it.valid && !((it.nlmask >> k) & 1u)) {
      u32 c = (cls >> (2 * k)) & 3u;
      if (c == ING_LEAD) c = lead_cls;
      if (c == ING_KEEP) keep |= 1u << k;
    }
  }
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
  __shared__ u8 sout[ING_CHUNK];
  u32 lp = before + incl - cnt;
#pragma unroll
  for (int k = 0; k < ING_PER; k++)
    if ((keep >> k) & 1u) sout[lp++] = (u8)ingest_code(it.b[k]);
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

// codes_dev needs room for nbytes bytes.  lead / rest: nchunks u32, last_ls: nchunks u64, hflag: nchunks bytes,
// partials: scan_partials_len(nchunks), total_dev: one u64.  Everything is enqueued on st; *total_dev holds n afterwards.
void ingest(const u8* raw_dev, u64 nbytes, u8* codes_dev, u32* lead, u32* rest, u64* last_ls, u8* hflag, u32* partials,
            u64* total_dev, bool fasta, cudaStream_t st) {
  const u64 nch = ingest_chunks(nbytes);
  if (!nch) return;
  k_ingest_summary<<<(unsigned)nch, ING_THREADS, 0, st>>>(raw_dev, nbytes, fasta, lead, rest, last_ls);
  CK_LAUNCH();
  k_ingest_carry<<<1, 1024, 0, st>>>(raw_dev, fasta, nch, lead, rest, last_ls, hflag, total_dev);
  CK_LAUNCH();
  scan_excl_u32(rest, nch, partials, nullptr, st);
  k_ingest_emit<<<(unsigned)nch, ING_THREADS, 0, st>>>(raw_dev, nbytes, fasta, rest, hflag, 0ull, codes_dev);
  CK_LAUNCH();
  g_launches += 3;
}

}  // namespace kiss
