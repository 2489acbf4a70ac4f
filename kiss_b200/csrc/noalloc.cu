// noalloc.cu — kiss_b200_ingest_noalloc_dev: the device ingest (ingest.cu) for callers that cannot allocate.
//
// No counterpart in the reference.  Some callers cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  kiss_b200_ingest_dev takes its
// per-chunk scratch (two kept-byte counts, the header-state function, the header flag, the scan partials, the total)
// from the context's arena, which may grow through cudaMalloc / cudaMallocAsync.  This variant runs the SAME kernels
// with that scratch in a file-scope __device__ array of fixed size: it makes no cudaMalloc, cudaFree or other
// allocating call, and in exchange accepts at most KISS_B200_NOALLOC_MAX_BYTES of file per call.
//
// The array exists once per GPU that loads the library (about 0.66 MB each, held until the process exits) and is shared
// by every call on that GPU: it keeps the last call's contents (every word a call reads is written by that call first —
// the grand total is zeroed by k_ingest_bscan — so nothing needs clearing between calls), and two calls that overlap on
// one GPU (two contexts, two host threads) race on it.
//
// In host code the array's NAME is a host-side stand-in, not a device address (a kernel handed it faults or silently
// uses the wrong memory): the device address comes from cudaGetSymbolAddress, once per device, with that device current.
#include <cstdio>
#include <string>

#include "../../include/kiss_b200.h"
#include "kernels.h"

using namespace kiss;

namespace {

constexpr u64 MAX_BYTES = KISS_B200_NOALLOC_MAX_BYTES;
constexpr u64 ING_CHUNK_BYTES = 4096;                       // ingest.cu: ING_THREADS * ING_PER
constexpr u64 MAX_CHUNKS = MAX_BYTES / ING_CHUNK_BYTES;     // 65536
constexpr u64 al256(u64 x) { return (x + 255) & ~255ull; }
// section offsets (bytes) inside the scratch array
constexpr u64 OFF_KEPT0 = 0;
constexpr u64 OFF_KEPT1 = OFF_KEPT0 + al256(MAX_CHUNKS * 4);
constexpr u64 OFF_FN = OFF_KEPT1 + al256(MAX_CHUNKS * 4);
constexpr u64 OFF_HFLAG = OFF_FN + al256(MAX_CHUNKS);
constexpr u64 OFF_BFN = OFF_HFLAG + al256(MAX_CHUNKS);
constexpr u64 OFF_PARTIALS = OFF_BFN + al256(MAX_CHUNKS / 1024 + 2);
constexpr u64 OFF_TOTAL = OFF_PARTIALS + al256((MAX_CHUNKS / 4096 + 2) * 4);
constexpr u64 SCRATCH_BYTES = OFF_TOTAL + 256;

}  // namespace

static __device__ __align__(256) unsigned char g_noalloc_scratch[SCRATCH_BYTES];

namespace {

constexpr int MAX_DEVICES = 64;
unsigned char* g_scratch_ptr[MAX_DEVICES] = {};  // device address of g_noalloc_scratch on each GPU, resolved on first use

// device address of this GPU's copy of the scratch array (`device` is current)
unsigned char* scratch_of(int device) {
  if (device < 0 || device >= MAX_DEVICES) throw CudaError(KISS_B200_ERR_ARG, "device index out of range");
  if (!g_scratch_ptr[device]) {
    void* p = nullptr;
    CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    g_scratch_ptr[device] = static_cast<unsigned char*>(p);
  }
  return g_scratch_ptr[device];
}

}  // namespace

extern "C" int kiss_b200_ingest_noalloc_dev(kiss_b200_ctx* c, const uint8_t* file_dev, uint64_t nbytes, int fasta,
                                            uint8_t* codes_dev, uint64_t* n_out) {
  if (!c || !n_out || (nbytes && (!file_dev || !codes_dev))) {
    api_set_last_error("bad argument");
    return KISS_B200_ERR_ARG;
  }
  *n_out = 0;
  if (nbytes > MAX_BYTES) {
    api_set_last_error("kiss_b200_ingest_noalloc_dev: " + std::to_string(nbytes) + " bytes exceed the fixed scratch of this entry (KISS_B200_NOALLOC_MAX_BYTES = " +
                       std::to_string(MAX_BYTES) + "); use kiss_b200_ingest_dev");
    return KISS_B200_ERR_TOO_LARGE;
  }
  if (nbytes == 0) return KISS_B200_OK;
  try {
    const int device = api_ctx_device(c);
    CK(cudaSetDevice(device));
    cudaStream_t st = static_cast<cudaStream_t>(kiss_b200_stream(c));
    unsigned char* s = scratch_of(device);
    u64* total = reinterpret_cast<u64*>(s + OFF_TOTAL);
    g_prof = nullptr;  // (the per-kernel profiler creates events on demand: not on this path)
    ingest(file_dev, nbytes, codes_dev, reinterpret_cast<u32*>(s + OFF_KEPT0), reinterpret_cast<u32*>(s + OFF_KEPT1),
           s + OFF_FN, s + OFF_BFN, s + OFF_HFLAG, reinterpret_cast<u32*>(s + OFF_PARTIALS), total, fasta != 0, st);
    u64 n = 0;
    CK(cudaMemcpyAsync(&n, total, sizeof(u64), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    *n_out = n;
    return KISS_B200_OK;
  } catch (const CudaError& e) {
    api_set_last_error(e.what());
    cudaGetLastError();
    return e.code;
  }
}
