// Drop-in for the reference's kiss2_sorter.hpp:7-50 (PREFIX_DOUBLING) — GPU backed, see kiss_b200_backend.hpp.
#pragma once
#include "kiss_b200_backend.hpp"

namespace biovoltron {
template<typename size_type = std::uint32_t>
struct KISS2Sorter : detail::B200Sorter<size_type, KISS_B200_PREFIX_DOUBLING> {};
}  // namespace biovoltron
