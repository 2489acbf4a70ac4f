// Drop-in for the reference's kiss1_sorter.hpp:7-50 (PARALLEL_SORTING) — GPU backed, see kiss_b200_backend.hpp.
// The general-alphabet get_suffix_array overloads are out of scope (never reached from the CLI, suffix_sort.hpp:58-60).
#pragma once
#include "kiss_b200_backend.hpp"

namespace biovoltron {
template<typename size_type = std::uint32_t>
struct KISS1Sorter : detail::B200Sorter<size_type, KISS_B200_PARALLEL_SORTING> {};
}  // namespace biovoltron
