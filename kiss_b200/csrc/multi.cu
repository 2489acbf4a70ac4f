// multi.cu — kiss_b200_suffix_sort_dna_u32_multi: the sharded LMS sort behind the C ABI, one process,
// one host thread and one context per GPU.
//
// Same partitioning as kiss_b200/multi_gpu.py (which runs it one PROCESS per GPU over torch.distributed for bench.py):
// rank r uploads slice r of the text over its own PCIe link and the GPUs complete it among themselves over NVLink; every
// GPU packs and classifies the text; rank r owns slice r of the text-ordered LMS list; splitters come from sampled
// first-16-base keys; ONE personalised exchange moves LMS positions (4 bytes each) — here as peer copies over NVLink /
// NVSwitch (cudaMemcpyPeerAsync): no NCCL is needed inside one process; every GPU sorts its key range with the
// single-GPU kernels and fetches the induction contexts of its suffixes; the (position, context) ranges are concatenated
// on rank 0 by peer copies; rank 0 places and induces (a single dependency chain: it does not shard); the finished array
// leaves in slices through every GPU's PCIe link.  Equal first keys share a rank, so tie rules stay local and the result
// is bit-identical to the single-GPU path.  The steps are the public device-stage entries of include/kiss_b200.h.
// PREFIX_DOUBLING and the unbounded order are not sharded (rank doubling is one global computation): they run on ctxs[0].
#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/kiss_b200.h"
#include "kernels.h"

using namespace kiss;

namespace {

struct Barrier {  // reusable; every thread always arrives (a failed rank keeps arriving and does nothing in between)
  std::mutex m;
  std::condition_variable cv;
  int n, waiting = 0, phase = 0;
  explicit Barrier(int n_) : n(n_) {}
  void wait() {
    std::unique_lock<std::mutex> lk(m);
    const int ph = phase;
    if (++waiting == n) { waiting = 0; phase++; cv.notify_all(); }
    else cv.wait(lk, [&] { return phase != ph; });
  }
};

struct Shared {
  int G;
  std::vector<std::vector<u32>> samples;     // per rank
  std::vector<u32> splitters;
  std::vector<std::vector<uint64_t>> counts; // counts[src][dst]
  std::vector<u32*> recv;                    // per rank: received positions
  std::vector<u64> recv_total;
  u32 *pos0 = nullptr, *ctx0 = nullptr;      // rank 0: the concatenated sorted order and its contexts
  std::vector<u8*> dS;                       // per rank: the text on its GPU
  u32* dSA0 = nullptr;                       // rank 0: the finished suffix array
  std::vector<int> device;
  std::mutex err_m;
  int err = 0;
  std::string err_msg;
  void fail(int code, const std::string& msg) {
    std::lock_guard<std::mutex> lk(err_m);
    if (!err) { err = code; err_msg = msg; }
  }
  bool ok() { std::lock_guard<std::mutex> lk(err_m); return err == 0; }
};

#define SAMPLES_PER_RANK 4096u

double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

void rank_main(int r, Shared& sh, Barrier& bar, kiss_b200_ctx** ctxs, const uint8_t* S_host, uint64_t n, uint32_t k,
               uint32_t* SA_host, kiss_b200_timings* t_out) {
  const int G = sh.G;
  kiss_b200_ctx* c = ctxs[r];
  u8* dS = nullptr;
  u32 *send = nullptr, *ctxbuf = nullptr, *dSA = nullptr, *slice = nullptr;
  // slice r of the text (16-byte granules) and of the suffix array: what rank r moves over its own PCIe link
  const uint64_t tchunk = ((n + G - 1) / G + 15) / 16 * 16, schunk = (n + 1 + G - 1) / G;
  const uint64_t tlo = std::min<uint64_t>(n, r * tchunk), thi = std::min<uint64_t>(n, (r + 1) * tchunk);
  const uint64_t slo = std::min<uint64_t>(n + 1, r * schunk), shi = std::min<uint64_t>(n + 1, (r + 1) * schunk);
  uint64_t m = 0, n_long = 0, lo = 0, hi = 0, total = 0;
  const uint32_t *keys = nullptr, *pos = nullptr;
  cudaStream_t st = nullptr;
  double t0 = now_s(), t_h2d = 0, t_prep = 0, t_sort = 0;
  // every step: skipped once any rank has failed; the barriers below are always reached
  auto step = [&](auto&& f) {
    if (!sh.ok()) return;
    try {
      f();
    } catch (const CudaError& e) {
      sh.fail(e.code, e.what());
    } catch (const std::exception& e) {
      sh.fail(KISS_B200_ERR_CUDA, e.what());
    }
  };
  auto api = [&](int rc) { if (rc != KISS_B200_OK) throw CudaError(rc, kiss_b200_last_error()); };

  step([&] {
    CK(cudaSetDevice(sh.device[r]));
    st = static_cast<cudaStream_t>(kiss_b200_stream(c));
    for (int p = 0; p < G; p++)
      if (p != r) { cudaDeviceEnablePeerAccess(sh.device[p], 0); cudaGetLastError(); }  // (already enabled / unsupported: staged copies)
    CK(cudaMalloc(&dS, n + 64));
    sh.dS[r] = dS;
    // 1 / G of the text through this GPU's own PCIe link ...
    if (thi > tlo) CK(cudaMemcpy(dS + tlo, S_host + tlo, thi - tlo, cudaMemcpyHostToDevice));
  });
  bar.wait();
  step([&] {
    // ... and the other slices from the GPUs that uploaded them, over NVLink (an all-gather by peer copies)
    for (int p = 0; p < G; p++) {
      const uint64_t plo = std::min<uint64_t>(n, p * tchunk), phi = std::min<uint64_t>(n, (p + 1) * tchunk);
      if (p != r && phi > plo) CK(cudaMemcpyPeerAsync(dS + plo, sh.device[r], sh.dS[p] + plo, sh.device[p], phi - plo, st));
    }
    CK(cudaStreamSynchronize(st));
    t_h2d = now_s() - t0;
    api(kiss_b200_prepare_dev(c, dS, n, k, &m));
    api(kiss_b200_lms_keys_dev(c, &keys, &pos, &n_long));
    t_prep = now_s() - t0 - t_h2d;
    lo = n_long * (uint64_t)r / G;
    hi = n_long * (uint64_t)(r + 1) / G;
    // strided sample of this slice's keys
    const uint64_t cnt = std::min<uint64_t>(SAMPLES_PER_RANK, hi - lo);
    sh.samples[r].resize(cnt);
    if (cnt) {
      const uint64_t stride = (hi - lo) / cnt;
      CK(cudaMemcpy2D(sh.samples[r].data(), sizeof(u32), keys + lo, stride * sizeof(u32), sizeof(u32), cnt, cudaMemcpyDeviceToHost));
    }
  });
  bar.wait();
  if (r == 0)
    step([&] {
      std::vector<u32> all;
      for (auto& v : sh.samples) all.insert(all.end(), v.begin(), v.end());
      std::sort(all.begin(), all.end());
      sh.splitters.clear();
      if (!all.empty())
        for (int d = 1; d < G; d++) sh.splitters.push_back(all[(uint64_t)d * all.size() / G]);  // destination = # splitters <= key
      else
        sh.splitters.assign(G - 1, 0xFFFFFFFFu);
    });
  bar.wait();
  step([&] {
    CK(cudaMalloc(&send, std::max<uint64_t>(hi - lo, 1) * sizeof(u32)));
    sh.counts[r].assign(G, 0);
    api(kiss_b200_partition_lms_dev(c, sh.splitters.data(), G - 1, lo, hi, send, sh.counts[r].data()));
  });
  bar.wait();
  step([&] {
    total = 0;
    for (int s = 0; s < G; s++) total += sh.counts[s][r];
    sh.recv_total[r] = total;
    CK(cudaMalloc(&sh.recv[r], std::max<uint64_t>(total, 1) * sizeof(u32)));
    if (r == 0) {
      CK(cudaMalloc(&sh.pos0, std::max<uint64_t>(n_long, 4) * sizeof(u32)));
      CK(cudaMalloc(&sh.ctx0, std::max<uint64_t>(n_long, 4) * sizeof(u32)));
    }
  });
  bar.wait();
  step([&] {  // the exchange: this rank's group for rank d goes to d's buffer after the groups of the ranks before this one
    uint64_t soff = 0;
    for (int d = 0; d < G; d++) {
      const uint64_t cnt = sh.counts[r][d];
      uint64_t doff = 0;
      for (int s = 0; s < r; s++) doff += sh.counts[s][d];
      if (cnt) CK(cudaMemcpyPeerAsync(sh.recv[d] + doff, sh.device[d], send + soff, sh.device[r], cnt * sizeof(u32), st));
      soff += cnt;
    }
    CK(cudaStreamSynchronize(st));
  });
  bar.wait();
  step([&] {  // chunks arrived in source-rank order = ascending text order: the stable sort's precondition
    const double s0 = now_s();
    api(kiss_b200_sort_lms_subset_dev(c, sh.recv[r], total, nullptr));
    CK(cudaMalloc(&ctxbuf, std::max<uint64_t>(total, 1) * sizeof(u32)));
    api(kiss_b200_lms_contexts_dev(c, sh.recv[r], total, ctxbuf));
    uint64_t off = 0;
    for (int s = 0; s < r; s++) off += sh.recv_total[s];
    if (total) {
      CK(cudaMemcpyPeerAsync(sh.pos0 + off, sh.device[0], sh.recv[r], sh.device[r], total * sizeof(u32), st));
      CK(cudaMemcpyPeerAsync(sh.ctx0 + off, sh.device[0], ctxbuf, sh.device[r], total * sizeof(u32), st));
    }
    CK(cudaStreamSynchronize(st));
    t_sort = now_s() - s0;
  });
  bar.wait();
  if (r == 0)
    step([&] {
      CK(cudaMalloc(&dSA, (n + 1) * sizeof(u32)));
      kiss_b200_timings ft;
      api(kiss_b200_finish_from_sorted_ctx_dev(c, sh.pos0, sh.ctx0, n_long, dSA, &ft));
      sh.dSA0 = dSA;
      if (t_out) {
        *t_out = ft;
        t_out->h2d = t_h2d; t_out->get_lms = t_prep; t_out->lms_sort = t_sort;
        t_out->n = n; t_out->m = m;
        t_out->total_device = t_out->get_lms + t_out->lms_sort + t_out->put_lms + t_out->induce;
      }
    });
  bar.wait();
  // the array leaves through every GPU's PCIe link: slice r crosses NVLink to GPU r and goes to the host from there
  const double d0 = now_s();
  step([&] {
    if (shi > slo) {
      const u32* src = sh.dSA0 + slo;
      if (r != 0) {
        CK(cudaMalloc(&slice, (shi - slo) * sizeof(u32)));
        CK(cudaMemcpyPeerAsync(slice, sh.device[r], sh.dSA0 + slo, sh.device[0], (shi - slo) * sizeof(u32), st));
        CK(cudaStreamSynchronize(st));
        src = slice;
      }
      CK(cudaMemcpy(SA_host + slo, src, (shi - slo) * sizeof(u32), cudaMemcpyDeviceToHost));
    }
  });
  bar.wait();
  if (r == 0 && t_out) t_out->d2h = now_s() - d0;
  // release (peers are done with this rank's buffers after the last barrier)
  cudaSetDevice(sh.device[r]);
  if (dS) cudaFree(dS);
  if (send) cudaFree(send);
  if (ctxbuf) cudaFree(ctxbuf);
  if (slice) cudaFree(slice);
  if (sh.recv[r]) cudaFree(sh.recv[r]);
  if (r == 0) {
    if (sh.pos0) cudaFree(sh.pos0);
    if (sh.ctx0) cudaFree(sh.ctx0);
    if (dSA) cudaFree(dSA);
  }
}

}  // namespace

extern "C" int kiss_b200_suffix_sort_dna_u32_multi(kiss_b200_ctx** ctxs, int n_gpus, const uint8_t* S_host, uint64_t n,
                                                   uint32_t k, int algo, uint32_t* SA_host, kiss_b200_timings* t) {
  if (!ctxs || n_gpus < 1 || n_gpus > 64 || !SA_host || (n && !S_host) || (algo != 0 && algo != 1)) {
    api_set_last_error("bad argument");
    return KISS_B200_ERR_ARG;
  }
  for (int r = 0; r < n_gpus; r++)
    if (!ctxs[r]) { api_set_last_error("null context"); return KISS_B200_ERR_ARG; }
  // one GPU, tiny texts, and the orders that are one global computation: the single-GPU entry on ctxs[0]
  if (n_gpus == 1 || n < 65536 || algo == KISS_B200_PREFIX_DOUBLING || (uint64_t)k + 1 >= n)
    return kiss_b200_suffix_sort_dna_u32(ctxs[0], S_host, n, k, algo, SA_host, t);
  if (n + 1 >= 0xFFFFFFFFull) { api_set_last_error("text too long for 32-bit suffix indices"); return KISS_B200_ERR_TOO_LARGE; }
  Shared sh;
  sh.G = n_gpus;
  sh.samples.resize(n_gpus); sh.counts.resize(n_gpus); sh.dS.assign(n_gpus, nullptr); sh.recv.assign(n_gpus, nullptr); sh.recv_total.assign(n_gpus, 0);
  for (int r = 0; r < n_gpus; r++) sh.device.push_back(api_ctx_device(ctxs[r]));
  Barrier bar(n_gpus);
  std::vector<std::thread> th;
  for (int r = 1; r < n_gpus; r++) th.emplace_back(rank_main, r, std::ref(sh), std::ref(bar), ctxs, S_host, n, k, SA_host, t);
  rank_main(0, sh, bar, ctxs, S_host, n, k, SA_host, t);
  for (auto& x : th) x.join();
  if (sh.err) { api_set_last_error(sh.err_msg); return sh.err; }
  return KISS_B200_OK;
}
