// oracle/ref_harness.cpp — TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// A thin extern "C" harness around the *unmodified* jhhung/kISS headers, compiled
// from where they lie under /root/reference by oracle/build_ref.sh into
// oracle/_ref/libkiss_ref.so.  It exposes the reference's own suffix_sort path
// (kiss1_suffix_array_dna: kiss1_core.hpp:229-268, kiss2_suffix_array_dna:
// kiss2_core.hpp:888-896) plus a stage-by-stage replay of the kiss1 driver that
// snapshots the intermediate arrays, so that the C restatement (oracle/kiss_oracle.c)
// and the CUDA kernels can be pinned stage by stage.
//
// Nothing in here restates the algorithm: every stage is a call into the
// reference's function templates.
#define SPDLOG_ACTIVE_LEVEL SPDLOG_LEVEL_OFF
#include <biovoltron/algo/sort/kiss1_sorter.hpp>
#include <biovoltron/algo/sort/kiss2_sorter.hpp>
#include <biovoltron/algo/align/exact_match/fm_index.hpp>
#include <biovoltron/file_io/fasta.hpp>
#include <utils/io.hpp>
#include <fstream>
#include <sstream>

#include <chrono>
#include <cstdint>
#include <cstring>

using biovoltron::kiss::vector;
namespace kiss = biovoltron::kiss;

static double now_s() {
  using namespace std::chrono;
  return duration<double>(steady_clock::now().time_since_epoch()).count();
}

extern "C" {

// Full reference path, PARALLEL_SORTING.  SA_out has n+1 slots.  Returns seconds spent in
// kiss1_suffix_array_dna (the region suffix_sort.hpp:57-61 / experiment/src/kiss-1.cpp:42-47 time).
double ref_kiss1_suffix_array_dna(const uint8_t* S_, uint64_t n, uint32_t k, uint64_t threads,
                                  uint32_t* SA_out) {
  vector<uint8_t> S(S_, S_ + n);
  vector<uint32_t> SA;
  double t0 = now_s();
  kiss::kiss1_suffix_array_dna<uint8_t, uint32_t>(S, SA, k, threads);
  double t1 = now_s();
  if (SA_out) std::memcpy(SA_out, SA.data(), sizeof(uint32_t) * SA.size());
  return t1 - t0;
}

// Full reference path, PREFIX_DOUBLING.
double ref_kiss2_suffix_array_dna(const uint8_t* S_, uint64_t n, uint32_t k, uint64_t threads,
                                  uint32_t* SA_out) {
  vector<uint8_t> S(S_, S_ + n);
  vector<uint32_t> SA;
  double t0 = now_s();
  kiss::kiss2_suffix_array_dna<uint8_t, uint32_t>(S, SA, k, threads);
  double t1 = now_s();
  if (SA_out) std::memcpy(SA_out, SA.data(), sizeof(uint32_t) * (n + 1));
  return t1 - t0;
}

// General alphabet (kiss1_core.hpp:270-311 / kiss2_core.hpp:898-906): S_ holds n bytes of any values.
double ref_kiss1_suffix_array(const uint8_t* S_, uint64_t n, uint32_t k, uint64_t threads, uint32_t* SA_out) {
  vector<uint8_t> S(S_, S_ + n);
  vector<uint32_t> SA;
  double t0 = now_s();
  kiss::kiss1_suffix_array<uint8_t, uint32_t>(S, SA, k, threads);
  double t1 = now_s();
  if (SA_out) std::memcpy(SA_out, SA.data(), sizeof(uint32_t) * (n + 1));
  return t1 - t0;
}
double ref_kiss2_suffix_array(const uint8_t* S_, uint64_t n, uint32_t k, uint64_t threads, uint32_t* SA_out) {
  vector<uint8_t> S(S_, S_ + n);
  vector<uint32_t> SA;
  double t0 = now_s();
  kiss::kiss2_suffix_array<uint8_t, uint32_t>(S, SA, k, threads);
  double t1 = now_s();
  if (SA_out) std::memcpy(SA_out, SA.data(), sizeof(uint32_t) * (n + 1));
  return t1 - t0;
}

// kISS-2's LMS sort alone: the steps of kiss2_suffix_array_internal (kiss2_core.hpp:835-886) up to and including
// prefix_doubling (:874), same calls in the same order.  sorted_out receives the m LMS suffixes in the order
// place_back_lms leaves them (:690-725; sorted_out[0] == n).  Returns m.  n must be > 0.
uint32_t ref_kiss2_sorted_lms(const uint8_t* S_, uint64_t n_, uint32_t k, uint64_t threads, uint32_t* sorted_out) {
  vector<uint8_t> S(S_, S_ + n_);
  vector<uint32_t> SA;
  uint32_t n = (uint32_t)S.size();
  SA.reserve((size_t)n + 1);
  auto states = vector<kiss::ThreadState<uint32_t>>(threads);
  SA.resize((size_t)n + 1);
  auto m = kiss::get_lms(S, SA, states);
  const uint32_t bits_for_character = std::bit_width(4u - 1u);
  const uint32_t l = 8 * sizeof(uint32_t) / bits_for_character;
  auto m_encoded = kiss::get_encoded_S_length(SA, m, l, threads, states);
  if (uint64_t(m_encoded) * 3 > (uint64_t)n + 1) SA.resize(uint64_t(m_encoded) * 3);
  kiss::prefix_doubling<uint8_t, uint32_t>(S, SA, k, bits_for_character, l, m, m_encoded, threads, states);
  if (sorted_out) std::memcpy(sorted_out, SA.data(), sizeof(uint32_t) * m);
  return m;
}

// Stage replay of kiss1_suffix_array_dna (kiss1_core.hpp:229-268), same calls in the same
// order, with snapshots.  Any out pointer may be NULL.
//   lms_out      : m entries  — get_lms output (ascending positions, sentinel n last)
//   sorted_out   : m entries  — after lms_suffix_direct_sort_dna (sorted_out[0] == n)
//   placed_out   : n+1        — after put_lms_suffix (EMPTY = 0xFFFFFFFF elsewhere)
//   sa_out       : n+1        — after induced_sort
//   counters_out : 5*256 u32  — sum over threads of ThreadState::bucket (kiss_common.hpp:519,529-536)
//   stage_sec    : 5 doubles  — resize, get_lms, lms sort, put, induce
// Returns m (LMS count including the sentinel).  n must be > 0.
uint32_t ref_kiss1_stages(const uint8_t* S_, uint64_t n_, uint32_t k, uint64_t threads,
                          uint32_t* lms_out, uint32_t* sorted_out, uint32_t* placed_out,
                          uint32_t* sa_out, uint32_t* counters_out, double* stage_sec) {
  vector<uint8_t> S(S_, S_ + n_);
  vector<uint32_t> SA;
  uint32_t n = (uint32_t)S.size();
  SA.reserve((size_t)n + 1);
  auto states = vector<kiss::ThreadState<uint32_t>>(threads);
  double t[6];
  t[0] = now_s();
  SA.resize((size_t)n + 1);
  t[1] = now_s();
  auto m = kiss::get_lms(S, SA, states);
  t[2] = now_s();
  if (lms_out) std::memcpy(lms_out, SA.data(), sizeof(uint32_t) * m);
  if (counters_out) {
    std::memset(counters_out, 0, sizeof(uint32_t) * 5 * 256);
    for (auto& st : states)
      for (size_t i = 0; i < 5 * 256; i++) counters_out[i] += st.bucket[i];
  }
  double t2b = now_s();
  auto buffer = std::ranges::subrange(std::begin(SA) + (n / 2), std::end(SA));
  auto output_SA = std::ranges::subrange(std::begin(SA), std::begin(SA) + (n / 2));
  kiss::lms_suffix_direct_sort_dna<uint8_t, uint32_t>(S, output_SA, buffer, m, k, threads, states);
  t[3] = now_s();
  if (sorted_out) std::memcpy(sorted_out, SA.data(), sizeof(uint32_t) * m);
  double t3b = now_s();
  auto SA1 = std::ranges::subrange(SA.begin() + 1, SA.begin() + m);
  kiss::put_lms_suffix(S, SA, SA1, states);
  t[4] = now_s();
  if (placed_out) std::memcpy(placed_out, SA.data(), sizeof(uint32_t) * ((size_t)n + 1));
  double t4b = now_s();
  kiss::induced_sort(S, SA, states);
  t[5] = now_s();
  if (sa_out) std::memcpy(sa_out, SA.data(), sizeof(uint32_t) * ((size_t)n + 1));
  if (stage_sec) {
    stage_sec[0] = t[1] - t[0];
    stage_sec[1] = t[2] - t[1];
    stage_sec[2] = t[3] - t2b;
    stage_sec[3] = t[4] - t3b;
    stage_sec[4] = t[5] - t4b;
  }
  return m;
}

// put_lms_suffix + induced_sort of the reference applied to a caller-supplied sorted LMS
// list (sorted[0] must be n).  Used to validate an LMS order model independently of the
// reference's comparator.
void ref_induce_from_sorted_lms(const uint8_t* S_, uint64_t n_, uint64_t threads,
                                const uint32_t* sorted, uint32_t* sa_out) {
  vector<uint8_t> S(S_, S_ + n_);
  vector<uint32_t> SA;
  uint32_t n = (uint32_t)S.size();
  auto states = vector<kiss::ThreadState<uint32_t>>(threads);
  SA.resize((size_t)n + 1);
  auto m = kiss::get_lms(S, SA, states);
  std::memcpy(SA.data(), sorted, sizeof(uint32_t) * m);
  auto SA1 = std::ranges::subrange(SA.begin() + 1, SA.begin() + m);
  kiss::put_lms_suffix(S, SA, SA1, states);
  kiss::induced_sort(S, SA, states);
  std::memcpy(sa_out, SA.data(), sizeof(uint32_t) * ((size_t)n + 1));
}

// the reference's own ingest: read_sequence (include/utils/io.hpp:6-18) on the bytes of a file, then the `% 4` of
// suffix_sort_main (include/command/suffix_sort.hpp:33).  Returns the text length; out needs room for nbytes codes.
uint64_t ref_read_sequence(const uint8_t* file, uint64_t nbytes, uint8_t* out) {
  std::istringstream fin(std::string(reinterpret_cast<const char*>(file), nbytes));
  auto seq = read_sequence(fin);
  uint64_t n = seq.size();
  for (uint64_t i = 0; i < n; i++) out[i] = (uint8_t)(seq[i] % 4);
  return n;
}

// The reference's fmindex_build command (include/command/fmindex_build.hpp:27-34) on a code text: FMIndex<4, uint32_t,
// KISS1Sorter<uint32_t>>{.LOOKUP_LEN = 0}, build() (fm_index.hpp:380-388: suffix array with k = 32, then build_occ /
// build_bwt / build_sa / build_lookup) and save() (:591-617) into the file `path`.  Returns the text length.
uint64_t ref_fmindex_build_file(const uint8_t* S_, uint64_t n, const char* path) {
  biovoltron::istring seq(reinterpret_cast<const biovoltron::ichar*>(S_), n);
  auto fmi = biovoltron::FMIndex<4, uint32_t, biovoltron::KISS1Sorter<uint32_t>>{.LOOKUP_LEN = 0};
  fmi.build(seq);
  auto fout = std::ofstream{path};
  fmi.save(fout);
  return n;
}

int ref_abi_version() { return 1; }

}  // extern "C"
