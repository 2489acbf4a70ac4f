// option_a_tu.cpp — INTEGRATION.md "Option A" as a compiled fact: the REFERENCE's own headers (fm_index.hpp, istring, the
// SASorter concept) and its own call sites, unchanged, built with kiss_b200/host/include placed before the reference's
// include/ on the include path, so that biovoltron::KISS1Sorter / KISS2Sorter resolve to the B200 facade.
// Built by tools/build_option_a.sh where /root/reference exists (the binary travels to the GPU box); run by
// tests/test_option_a.py.  The text is a fixed LCG stream so that the test can rebuild it.
#define SPDLOG_ACTIVE_LEVEL SPDLOG_LEVEL_OFF
#include <biovoltron/algo/align/exact_match/fm_index.hpp>
#include <biovoltron/algo/sort/kiss2_sorter.hpp>
#include <fstream>
#include <string>
#include <string_view>
#include <cstdio>
int main(int argc, char** argv) {
  // the reference's own call sites, unchanged: tests/kiss.cpp:25 (istring + KISS1Sorter), fmindex_build.hpp:27-34
  auto seq = biovoltron::istring{};
  unsigned x = 12345;
  for (int i = 0; i < 100000; i++) { x = x * 1664525u + 1013904223u; seq.push_back((x >> 24) & 3); }
  auto sa = biovoltron::KISS1Sorter<uint32_t>::get_suffix_array_dna(seq, 256, 24);
  auto sa2 = biovoltron::KISS2Sorter<uint32_t>::get_suffix_array_dna(seq, 256);
  bool ok = sa.size() == seq.size() + 1 && sa[0] == seq.size();
  for (size_t i = 1; i < sa.size() && ok; i++) ok = seq.substr(sa[i - 1], 256) <= seq.substr(sa[i], 256);
  // tests/kiss.cpp:53-65 ("kISS-1 general"): symbols 'A'..'D' through get_suffix_array on a string_view
  auto gen = std::string{};
  for (int i = 0; i < 120000; i++) { x = x * 1664525u + 1013904223u; gen += (char)(65 + ((x >> 24) & 3)); }
  const auto gen_sv = std::string_view{gen};
  auto sag = biovoltron::KISS1Sorter<uint32_t>::get_suffix_array(gen_sv, 256, 24);
  ok = ok && sag.size() == gen.size() + 1 && sag[0] == gen.size();
  for (size_t i = 1; i < sag.size() && ok; i++) ok = gen_sv.substr(sag[i - 1], 256) <= gen_sv.substr(sag[i], 256);
  auto fmi = biovoltron::FMIndex<4, uint32_t, biovoltron::KISS1Sorter<uint32_t>>{.LOOKUP_LEN = 0};
  fmi.build(seq);
  if (argc > 1) { auto f = std::ofstream{argv[1]}; fmi.save(f); }
  std::printf("%s %zu %zu\n", ok ? "ok" : "BAD", sa.size(), sa2.size());
  return ok ? 0 : 1;
}
