// The reference's SASorter concept (sorter.hpp:7-10): "t.get_suffix_array_dna(range) is valid".
#pragma once
#include <concepts>
#include <cstdint>
#include <vector>

namespace biovoltron {
template<typename T>
concept SASorter = requires(T t, const std::vector<std::uint8_t>& ref) { t.get_suffix_array_dna(ref); };
}  // namespace biovoltron
