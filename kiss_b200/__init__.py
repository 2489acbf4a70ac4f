"""kiss_b200 — B200-native k-ordered suffix-array construction (the `suffix_sort` path of jhhung/kISS).

The product is kiss_b200/libkiss_b200.so (hand-written sm_100a CUDA behind the C ABI of
include/kiss_b200.h).  This package is the thin Python host side used by the tests and the benchmark:
`KISS1Sorter` / `KISS2Sorter` mirror the reference's sorter facades
(include/biovoltron/algo/sort/kiss1_sorter.hpp:7-50, kiss2_sorter.hpp:7-50).
"""
from ._cabi import (EMPTY, PARALLEL_SORTING, PREFIX_DOUBLING, UNBOUNDED_K, KissB200Error, Timings, lib)
from .sorter import Context, KISS1Sorter, KISS2Sorter, default_context, suffix_array_dna_multi

__all__ = ["Context", "KISS1Sorter", "KISS2Sorter", "default_context", "suffix_array_dna_multi", "KissB200Error", "Timings", "lib",
           "EMPTY", "PARALLEL_SORTING", "PREFIX_DOUBLING", "UNBOUNDED_K"]
