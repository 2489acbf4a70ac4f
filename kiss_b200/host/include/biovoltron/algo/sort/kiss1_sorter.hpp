// Drop-in for the reference's kiss1_sorter.hpp:7-50 (PARALLEL_SORTING) — GPU backed, see kiss_b200_backend.hpp.
// get_suffix_array_dna and the general-alphabet get_suffix_array (kiss1_sorter.hpp:28-44) are both provided.
#pragma once
#include "kiss_b200_backend.hpp"

namespace biovoltron {
template<typename size_type = std::uint32_t>
struct KISS1Sorter : detail::B200Sorter<size_type, KISS_B200_PARALLEL_SORTING> {};
}  // namespace biovoltron
