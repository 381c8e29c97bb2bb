/* oracle_shim.c -- TEST INFRASTRUCTURE.  The three C-ABI symbols radix_sort.hpp
 * uses, answered by the CPU oracle (oracle/liboracle.so), so that the LOGIC of
 * tests_rs.cpp can be checked on a box without a GPU.  Never part of the
 * product; the GPU test links tests_rs.cpp against libb200sort.so instead. */
#include <stddef.h>

int oracle_sort0(void *data, size_t n, int kind);

int b200sort_keys_host(void *keys, size_t n, int kind) { return n ? oracle_sort0(keys, n, kind) : 0; }

int b200sort_records_host(void *recs, size_t n, size_t rec_bytes, size_t key_off, int kind) {
    (void)kind;
    if (key_off != 0) return 1; /* the oracle's pairs carry the key in field 0 */
    return n ? oracle_sort0(recs, n, rec_bytes == 8 ? 100 : 101) : 0;
}

const char *b200sort_last_error(void) { return "oracle shim"; }
