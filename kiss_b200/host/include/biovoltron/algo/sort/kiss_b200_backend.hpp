// kiss_b200_backend.hpp — C++ host side above the C ABI (include/kiss_b200.h).
//
// Gives the reference's sorter facades (kiss1_sorter.hpp:7-50, kiss2_sorter.hpp:7-50) a GPU implementation:
// `biovoltron::KISS1Sorter<uint32_t>` / `KISS2Sorter<uint32_t>` keep their names, static member functions,
// argument meaning and defaults, so call sites such as suffix_sort_main (include/command/suffix_sort.hpp:37-60),
// FMIndex::build (fm_index.hpp:380-388) and tests/kiss.cpp:25 compile unchanged against this directory when it
// is placed before the reference's include/ on the include path.  Errors: the reference throws std::bad_alloc
// (structs.hpp:32-38); here a non-zero status becomes std::bad_alloc (out of memory) or std::runtime_error.
// No CPU fallback: without a usable GPU every call throws.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <iterator>
#include <memory>
#include <new>
#include <ranges>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "kiss_b200.h"

namespace biovoltron {
namespace kiss {

// the reference's kiss::vector is a std::vector with a 64-byte aligned allocator (structs.hpp:11-58);
// alignment is irrelevant to a device pipeline, the element type and interface are what callers use
template<typename T>
using vector = std::vector<T>;

namespace b200 {

inline int& device_index() {
  static int dev = 0;
  return dev;
}

struct CtxDeleter {
  void operator()(kiss_b200_ctx* c) const { kiss_b200_destroy(c); }
};

// one context per host thread (a context is not thread-safe; separate contexts are)
inline kiss_b200_ctx* context() {
  thread_local std::unique_ptr<kiss_b200_ctx, CtxDeleter> ctx;
  if (!ctx) {
    kiss_b200_ctx* raw = nullptr;
    int rc = kiss_b200_create(&raw, device_index());
    if (rc != KISS_B200_OK) throw std::runtime_error(std::string("kiss_b200: ") + kiss_b200_last_error());
    ctx.reset(raw);
  }
  return ctx.get();
}

// number of GPUs the sorter facades use: the opt-in knob of SURVEY.md 8(b) — `gpus() = N`, or the environment variable
// KISS_B200_GPUS; default 1 (the reference has no such notion).  Devices device_index() .. device_index() + N - 1.
inline int& gpus() {
  static int g = [] { const char* e = std::getenv("KISS_B200_GPUS"); int v = e ? std::atoi(e) : 1; return v > 0 ? v : 1; }();
  return g;
}
inline std::vector<kiss_b200_ctx*>& contexts_multi(int n) {
  thread_local std::vector<std::unique_ptr<kiss_b200_ctx, CtxDeleter>> own;
  thread_local std::vector<kiss_b200_ctx*> raw;
  while ((int)own.size() < n) {
    kiss_b200_ctx* r = nullptr;
    int rc = kiss_b200_create(&r, device_index() + (int)own.size());
    if (rc != KISS_B200_OK) throw std::runtime_error(std::string("kiss_b200: ") + kiss_b200_last_error());
    own.emplace_back(r);
    raw.push_back(r);
  }
  return raw;
}

inline void raise(int rc) {
  if (rc == KISS_B200_OK) return;
  if (rc == KISS_B200_ERR_NOMEM) throw std::bad_alloc();
  throw std::runtime_error(std::string("kiss_b200: ") + kiss_b200_strerror(rc) + ": " + kiss_b200_last_error());
}

inline kiss_b200_timings& last_timings() {
  thread_local kiss_b200_timings t{};
  return t;
}

// kiss1_suffix_array_dna / kiss2_suffix_array_dna (kiss1_core.hpp:229-235, kiss2_core.hpp:888-896); `codes` = n bytes of
// any one-byte character type (the reference's callers pass uint8_t, int8_t or char: only the low two bits are read)
inline void suffix_array_dna(const void* codes, std::size_t n, vector<std::uint32_t>& SA, std::uint32_t k, int algo) {
  SA.resize(n + 1);
  if (gpus() > 1) {
    auto& cs = contexts_multi(gpus());
    raise(kiss_b200_suffix_sort_dna_u32_multi(cs.data(), gpus(), static_cast<const std::uint8_t*>(codes), n, k, algo,
                                              SA.data(), &last_timings()));
    return;
  }
  raise(kiss_b200_suffix_sort_dna_u32(context(), static_cast<const std::uint8_t*>(codes), n, k, algo, SA.data(),
                                      &last_timings()));
}
inline void suffix_array_dna(const vector<std::uint8_t>& S, vector<std::uint32_t>& SA, std::uint32_t k, int algo) {
  suffix_array_dna(S.data(), S.size(), SA, k, algo);
}

// kiss1_suffix_array / kiss2_suffix_array, the general-alphabet drivers (kiss1_core.hpp:270-311, kiss2_core.hpp:898-906):
// `S` = n bytes of any values (include/kiss_b200.h: kiss_b200_suffix_sort_u32 for the order among suffixes sharing k symbols)
inline void suffix_array(const void* S, std::size_t n, vector<std::uint32_t>& SA, std::uint32_t k, int algo) {
  SA.resize(n + 1);
  raise(kiss_b200_suffix_sort_u32(context(), static_cast<const std::uint8_t*>(S), n, k, algo, SA.data(), &last_timings()));
}

// fmindex_build (include/command/fmindex_build.hpp:27-34): the bytes of a FASTA / text file -> the bytes of the .fmi file
// FMIndex<sa_intv, uint32_t, Sorter>{.LOOKUP_LEN = 0}::save writes (fm_index.hpp:591-617); the suffix array stays on the
// device.  Returns the text length.
inline std::uint64_t fmindex_image_file(const std::vector<std::uint8_t>& file, std::vector<std::uint8_t>& image,
                                        std::uint32_t k = 32u, int algo = KISS_B200_PARALLEL_SORTING,
                                        std::uint32_t sa_intv = 4u, double* fm_seconds = nullptr) {
  image.resize(kiss_b200_fmindex_image_bytes(file.size(), sa_intv));
  std::uint64_t n = 0, bytes = 0;
  raise(kiss_b200_fmindex_build_file_u32(context(), file.data(), file.size(), k, algo, sa_intv, image.data(), image.size(),
                                         &bytes, &n, &last_timings(), fm_seconds));
  image.resize(bytes);
  return n;
}

// read_sequence (include/utils/io.hpp:6-18) + `% 4` + the sorter in one call: `file` holds the bytes of a FASTA or
// plain-text file, header / newline stripping and base encoding run on the device (kiss_b200_suffix_sort_file_u32).
// Returns the text length n; SA has n + 1 entries.
inline std::uint64_t suffix_array_file(const std::vector<std::uint8_t>& file, vector<std::uint32_t>& SA, std::uint32_t k,
                                       int algo) {
  SA.resize(file.size() + 1);
  std::uint64_t n = 0;
  raise(kiss_b200_suffix_sort_file_u32(context(), file.data(), file.size(), k, algo, SA.data(), SA.size(), &n,
                                       &last_timings()));
  SA.resize(n + 1);
  return n;
}

}  // namespace b200

template<typename char_type, typename size_type>
void kiss1_suffix_array_dna(const vector<char_type>& S, vector<size_type>& SA, size_type k, std::size_t /*num_threads*/) {
  static_assert(sizeof(char_type) == 1 && sizeof(size_type) == 4, "the B200 path implements <uint8_t, uint32_t>");
  b200::suffix_array_dna(S.data(), S.size(), SA, k, KISS_B200_PARALLEL_SORTING);
}

template<typename char_type, typename size_type>
void kiss2_suffix_array_dna(const vector<char_type>& S, vector<size_type>& SA, size_type k, std::size_t /*num_threads*/) {
  static_assert(sizeof(char_type) == 1 && sizeof(size_type) == 4, "the B200 path implements <uint8_t, uint32_t>");
  b200::suffix_array_dna(S.data(), S.size(), SA, k, KISS_B200_PREFIX_DOUBLING);
}

template<typename char_type, typename size_type>
void kiss1_suffix_array(const vector<char_type>& S, vector<size_type>& SA, size_type k, std::size_t /*num_threads*/) {
  static_assert(sizeof(char_type) == 1 && sizeof(size_type) == 4, "the B200 path implements <uint8_t, uint32_t>");
  b200::suffix_array(S.data(), S.size(), SA, k, KISS_B200_PARALLEL_SORTING);
}

template<typename char_type, typename size_type>
void kiss2_suffix_array(const vector<char_type>& S, vector<size_type>& SA, size_type k, std::size_t /*num_threads*/) {
  static_assert(sizeof(char_type) == 1 && sizeof(size_type) == 4, "the B200 path implements <uint8_t, uint32_t>");
  b200::suffix_array(S.data(), S.size(), SA, k, KISS_B200_PREFIX_DOUBLING);
}

}  // namespace kiss

namespace detail {
template<typename size_type, int ALGO>
struct B200Sorter {
  static_assert(sizeof(size_type) == 4, "only 32-bit suffix indices (the CLI's choice, suffix_sort.hpp:37)");
  using SA_t = kiss::vector<size_type>;

  static auto get_suffix_array_dna(const std::ranges::random_access_range auto& ref, size_type k = 256u,
                                   const std::size_t num_threads = std::thread::hardware_concurrency()) {
    auto S = kiss::vector<std::uint8_t>(std::ranges::begin(ref), std::ranges::end(ref));
    return get_suffix_array_dna(S, k, num_threads);
  }

  static auto get_suffix_array_dna(const kiss::vector<std::uint8_t>& S, size_type k = 256u,
                                   const std::size_t /*num_threads*/ = std::thread::hardware_concurrency()) {
    auto SA = SA_t{};
    kiss::b200::suffix_array_dna(S, SA, k, ALGO);
    return SA;
  }

  // general alphabet (kiss1_sorter.hpp:28-44)
  static auto get_suffix_array(const std::ranges::random_access_range auto& ref, size_type k = 256u,
                               const std::size_t num_threads = std::thread::hardware_concurrency()) {
    auto S = kiss::vector<std::uint8_t>(std::ranges::begin(ref), std::ranges::end(ref));
    return get_suffix_array(S, k, num_threads);
  }

  static auto get_suffix_array(const kiss::vector<std::uint8_t>& S, size_type k = 256u,
                               const std::size_t /*num_threads*/ = std::thread::hardware_concurrency()) {
    auto SA = SA_t{};
    kiss::b200::suffix_array(S.data(), S.size(), SA, k, ALGO);
    return SA;
  }

  static auto prepare_aligned_ref(const std::ranges::random_access_range auto& ref) {
    return kiss::vector<std::uint8_t>(std::ranges::begin(ref), std::ranges::end(ref));
  }
};
}  // namespace detail

}  // namespace biovoltron
