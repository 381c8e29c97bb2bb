"""radix_sort_b200 -- B200 (sm_100a) LSD radix sort behind the sort API of
jgrodzki/radix_sort (`slice.radix_sort()`, src/radix_sort/mod.rs:18-20,61-176).

The product is libb200sort.so (C ABI: include/b200sort.h; kernels: csrc/).
This package is the host-side mirror of the reference interface over that ABI.
"""
from ._ffi import B200SortError
from .api import (FLAG_NO_SKIP, debug_force_rerank, debug_reranked_tiles, debug_small_sort, debug_narrow_fill, release_cached_memory, ISIZE, NUMBER_OF_DIGITS, USIZE, Kind, counting_sort, key_bytes, kind_of,
                  launch_count, num_variants, radix_sort, radix_sort0, radix_sort1, radix_sort2, radix_sort3, radix_sort4,
                  radix_sort5, radix_sort_records, set_host_pipeline_min_bytes, set_variant, sort_timed, temp_bytes)

__all__ = [
    "B200SortError", "Kind", "USIZE", "ISIZE", "NUMBER_OF_DIGITS", "FLAG_NO_SKIP",
    "radix_sort", "radix_sort_records", "radix_sort0", "radix_sort1", "radix_sort2", "radix_sort3",
    "radix_sort4", "radix_sort5", "counting_sort", "temp_bytes", "key_bytes", "kind_of",
    "launch_count", "debug_force_rerank", "debug_reranked_tiles", "debug_small_sort", "debug_narrow_fill", "release_cached_memory", "set_variant", "num_variants", "sort_timed", "set_host_pipeline_min_bytes",
]
