/*
 * oracle.c -- CPU restatement of the jgrodzki/radix_sort hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * `--impl reference` leg may load this library, and only as the checker or as
 * the timed CPU baseline.  The product (libb200sort.so) never links or calls it.
 *
 * PARITY STATUS: "parity unpinned" in the strict sense -- the reference holds no
 * golden vectors (every test draws from an unseeded thread_rng, tests.rs:28,46,
 * 136,178) and no Rust toolchain exists in this environment, so the reference
 * itself cannot be run to produce fixtures.  What pins this oracle instead is
 * the property the reference's own tests assert (tests.rs:7-23,146,167,183):
 * output == std sort (integers), == sort_by(total_cmp) (floats), == stable
 * sort_by_key (tuples).  tests/test_oracle.py checks exactly that against an
 * independent numpy implementation, plus the derived known-answer vectors of
 * SURVEY.md section 8(c).
 *
 * What is restated (reference file:line):
 *   get_digit, unsigned           src/radix_sort/radix_digits.rs:7-53
 *   get_digit, signed             src/radix_sort/radix_digits.rs:55-101
 *   get_digit, f32 / f64          src/radix_sort/radix_digits.rs:103-124
 *   get_digit, (T,U) -> .0        src/radix_sort/radix_digits.rs:126-136
 *   radix_sort0 (sequential LSD)  src/radix_sort/mod.rs:182-212   -> oracle_sort0
 *   radix_sort  (production, native threads + 96-element write buffers)
 *                                 src/radix_sort/mod.rs:61-176    -> oracle_sort_mt
 *   radix_sort4 (rayon variant)   src/radix_sort/mod.rs:405-466   -> oracle_sort4_mt
 *   (T,U) of any shape, sequential src/radix_sort/radix_digits.rs:126-136 + mod.rs:182-212
 *                                                                 -> oracle_sort0_records
 *
 * Elements are moved bit-for-bit; the key transform lives only inside
 * get_digit, exactly as in the reference (radix_digits.rs:59,110-112).
 */
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef unsigned __int128 u128_t;
typedef __int128 i128_t;

/* kind ids: identical to include/b200sort.h so tests can share the enum. */
enum {
    K_U8 = 0, K_U16, K_U32, K_U64, K_U128,
    K_I8, K_I16, K_I32, K_I64, K_I128,
    K_F32, K_F64,
    /* record kinds: key-value pairs, key = field 0 (radix_digits.rs:126-136) */
    K_PAIR_U32 = 100, K_PAIR_U64 = 101
};

typedef struct { uint32_t k, v; } pair32_t;
typedef struct { uint64_t k, v; } pair64_t;

/* ---- get_digit, one inline function per built-in kind ------------------- */
/* unsigned: (*self >> index * 8) as u8              radix_digits.rs:11,19,27,35,43 */
static inline unsigned dg_u8(const uint8_t *e, unsigned i)  { return (uint8_t)(*e >> (i * 8)); }
static inline unsigned dg_u16(const uint16_t *e, unsigned i){ return (uint8_t)(*e >> (i * 8)); }
static inline unsigned dg_u32(const uint32_t *e, unsigned i){ return (uint8_t)(*e >> (i * 8)); }
static inline unsigned dg_u64(const uint64_t *e, unsigned i){ return (uint8_t)(*e >> (i * 8)); }
static inline unsigned dg_u128(const u128_t *e, unsigned i) { return (uint8_t)(*e >> (i * 8)); }
/* signed: ((*self ^ MIN) >> index * 8) as u8 -- arithmetic shift, low byte kept
 *                                                    radix_digits.rs:59,67,75,83,91 */
static inline unsigned dg_i8(const int8_t *e, unsigned i)   { return (uint8_t)(((int8_t)(*e ^ INT8_MIN)) >> (i * 8)); }
static inline unsigned dg_i16(const int16_t *e, unsigned i) { return (uint8_t)(((int16_t)(*e ^ INT16_MIN)) >> (i * 8)); }
static inline unsigned dg_i32(const int32_t *e, unsigned i) { return (uint8_t)((*e ^ INT32_MIN) >> (i * 8)); }
static inline unsigned dg_i64(const int64_t *e, unsigned i) { return (uint8_t)((*e ^ INT64_MIN) >> (i * 8)); }
static inline unsigned dg_i128(const i128_t *e, unsigned i) {
    const i128_t min128 = (i128_t)((u128_t)1 << 127);
    return (uint8_t)((*e ^ min128) >> (i * 8));
}
/* floats: b = to_bits() as iN; b ^= (b >> N-1) | iN::MIN; (b as uN >> index*8) as u8
 *                                                    radix_digits.rs:110-112,120-122 */
static inline unsigned dg_f32(const uint32_t *e, unsigned i) {
    int32_t b = (int32_t)*e;
    b ^= (b >> 31) | INT32_MIN;
    return (uint8_t)((uint32_t)b >> (i * 8));
}
static inline unsigned dg_f64(const uint64_t *e, unsigned i) {
    int64_t b = (int64_t)*e;
    b ^= (b >> 63) | INT64_MIN;
    return (uint8_t)((uint64_t)b >> (i * 8));
}
/* tuples: self.0.get_digit(index)                    radix_digits.rs:133-135 */
static inline unsigned dg_p32(const pair32_t *e, unsigned i) { return dg_u32(&e->k, i); }
static inline unsigned dg_p64(const pair64_t *e, unsigned i) { return dg_u64(&e->k, i); }

/* ---- radix_sort0: sequential LSD (mod.rs:182-212) ------------------------ */
/* ping-pong self<->temp (mod.rs:186-190), 256-bin histogram (:191-194),
 * in-place exclusive scan (:195-202), stable scatter (:203-207),
 * odd-digit copy back (:209-211). */
#define DEF_SORT0(NAME, T, ND, DG)                                            \
static int sort0_##NAME(T *data, size_t n) {                                  \
    if (n == 0) return 0;                                                     \
    T *temp = (T *)malloc(n * sizeof(T));                                     \
    if (!temp) return -1;                                                     \
    for (unsigned d = 0; d < (ND); ++d) {                                     \
        const T *src = (d % 2 == 0) ? data : temp;                            \
        T *dst = (d % 2 == 0) ? temp : data;                                  \
        size_t bins[256];                                                     \
        memset(bins, 0, sizeof bins);                                         \
        for (size_t i = 0; i < n; ++i) bins[DG(&src[i], d)] += 1;             \
        size_t start = 0;                                                     \
        for (int b = 0; b < 256; ++b) {                                       \
            size_t next = start + bins[b];                                    \
            bins[b] = start;                                                  \
            start = next;                                                     \
        }                                                                     \
        for (size_t i = 0; i < n; ++i) {                                      \
            unsigned g = DG(&src[i], d);                                      \
            memcpy(&dst[bins[g]], &src[i], sizeof(T));                        \
            bins[g] += 1;                                                     \
        }                                                                     \
    }                                                                         \
    if ((ND) % 2 == 1) memcpy(data, temp, n * sizeof(T));                     \
    free(temp);                                                               \
    return 0;                                                                 \
}

/* ---- radix_sort: production path (mod.rs:61-176) ------------------------- */
/* chunk = ceil(n / P) (mod.rs:66-70); uninitialised temp, one byte per 4 KiB
 * page touched serially by the caller (:71-82); per digit: one thread per chunk
 * builds a histogram (:90-109), serial digit-major / chunk-minor exclusive scan
 * (:110-120), one thread per chunk scatters through 256 x 96-element staging
 * buffers flushed by memcpy (:121-168); odd digit count copies back (:170-174).
 * Threads are created and joined twice per digit, as thread::scope does.
 * Deviation: n == 0 returns instead of panicking. */
#define BUFFER_SIZE 96
#define PAGE_SIZE 4096

typedef struct {
    const void *src; void *dst; size_t lo, hi; unsigned digit; size_t *bins;
} mt_job_t;

#define DEF_SORT_MT(NAME, T, ND, DG)                                          \
static void *mt_hist_##NAME(void *arg) {                                      \
    mt_job_t *j = (mt_job_t *)arg;                                            \
    const T *src = (const T *)j->src;                                         \
    size_t h[256];                                                            \
    memset(h, 0, sizeof h);                                                   \
    for (size_t i = j->lo; i < j->hi; ++i) h[DG(&src[i], j->digit)] += 1;     \
    memcpy(j->bins, h, sizeof h);                                             \
    return NULL;                                                              \
}                                                                             \
static void *mt_scatter_##NAME(void *arg) {                                   \
    mt_job_t *j = (mt_job_t *)arg;                                            \
    const T *src = (const T *)j->src;                                         \
    T *dst = (T *)j->dst;                                                     \
    size_t *starts = j->bins;                                                 \
    T (*buf)[BUFFER_SIZE] = (T (*)[BUFFER_SIZE])malloc(sizeof(T) * BUFFER_SIZE * 256); \
    size_t cnt[256];                                                          \
    memset(cnt, 0, sizeof cnt);                                               \
    for (size_t i = j->lo; i < j->hi; ++i) {                                  \
        unsigned g = DG(&src[i], j->digit);                                   \
        memcpy(&buf[g][cnt[g]], &src[i], sizeof(T));                          \
        cnt[g] += 1;                                                          \
        if (cnt[g] == BUFFER_SIZE) {                                          \
            memcpy(&dst[starts[g]], buf[g], sizeof(T) * BUFFER_SIZE);         \
            starts[g] += BUFFER_SIZE;                                         \
            cnt[g] = 0;                                                       \
        }                                                                     \
    }                                                                         \
    for (int g = 0; g < 256; ++g)                                             \
        if (cnt[g] > 0) memcpy(&dst[starts[g]], buf[g], sizeof(T) * cnt[g]);  \
    free(buf);                                                                \
    return NULL;                                                              \
}                                                                             \
static int sort_mt_##NAME(T *data, size_t n, int P) {                         \
    if (n == 0) return 0;                                                     \
    if (P < 1) P = 1;                                                         \
    size_t per = (n + (size_t)P - 1) / (size_t)P;                             \
    size_t nchunks = (n + per - 1) / per;                                     \
    T *temp = (T *)malloc(n * sizeof(T));                                     \
    if (!temp) return -1;                                                     \
    for (size_t b = 0; b < n * sizeof(T); b += PAGE_SIZE) ((volatile uint8_t *)temp)[b] = 0; \
    size_t *bins = (size_t *)malloc(nchunks * 256 * sizeof(size_t));          \
    mt_job_t *jobs = (mt_job_t *)malloc(nchunks * sizeof(mt_job_t));          \
    pthread_t *th = (pthread_t *)malloc(nchunks * sizeof(pthread_t));         \
    for (unsigned d = 0; d < (ND); ++d) {                                     \
        const T *src = (d % 2 == 0) ? data : temp;                            \
        T *dst = (d % 2 == 0) ? temp : data;                                  \
        for (size_t c = 0; c < nchunks; ++c) {                                \
            size_t lo = c * per, hi = lo + per < n ? lo + per : n;            \
            jobs[c] = (mt_job_t){src, dst, lo, hi, d, bins + c * 256};        \
            pthread_create(&th[c], NULL, mt_hist_##NAME, &jobs[c]);           \
        }                                                                     \
        for (size_t c = 0; c < nchunks; ++c) pthread_join(th[c], NULL);       \
        size_t prefix = 0;                                                    \
        for (int g = 0; g < 256; ++g)                                         \
            for (size_t c = 0; c < nchunks; ++c) {                            \
                size_t next = prefix + bins[c * 256 + g];                     \
                bins[c * 256 + g] = prefix;                                   \
                prefix = next;                                                \
            }                                                                 \
        for (size_t c = 0; c < nchunks; ++c)                                  \
            pthread_create(&th[c], NULL, mt_scatter_##NAME, &jobs[c]);        \
        for (size_t c = 0; c < nchunks; ++c) pthread_join(th[c], NULL);       \
    }                                                                         \
    if ((ND) % 2 == 1) memcpy(data, temp, n * sizeof(T));                     \
    free(th); free(jobs); free(bins); free(temp);                             \
    return 0;                                                                 \
}

/* ---- radix_sort4: the "rayon" path (mod.rs:405-466) ---------------------- */
/* chunk = ceil(n / (threads * CHUNK_MULTIPLIER)) (mod.rs:407-411); page-touched
 * uninitialised temp (:412-423); par_chunks histogram collected in chunk order
 * (:430-439); the same serial scan (:440-450); par_chunks direct, unbuffered
 * scatter (:451-461).  rayon's work-stealing pool is restated as an OpenMP
 * parallel-for with dynamic scheduling: a persistent pool whose idle workers
 * claim the next chunk.  Results do not depend on the schedule. */
#define CHUNK_MULTIPLIER 2
#define DEF_SORT4_MT(NAME, T, ND, DG)                                         \
static int sort4_mt_##NAME(T *data, size_t n, int P) {                        \
    if (n == 0) return 0;                                                     \
    if (P < 1) P = 1;                                                         \
    size_t per = (n + (size_t)P * CHUNK_MULTIPLIER - 1) / ((size_t)P * CHUNK_MULTIPLIER); \
    long nchunks = (long)((n + per - 1) / per);                               \
    T *temp = (T *)malloc(n * sizeof(T));                                     \
    if (!temp) return -1;                                                     \
    for (size_t b = 0; b < n * sizeof(T); b += PAGE_SIZE) ((volatile uint8_t *)temp)[b] = 0; \
    size_t *bins = (size_t *)malloc((size_t)nchunks * 256 * sizeof(size_t));  \
    for (unsigned d = 0; d < (ND); ++d) {                                     \
        const T *src = (d % 2 == 0) ? data : temp;                            \
        T *dst = (d % 2 == 0) ? temp : data;                                  \
        _Pragma("omp parallel for schedule(dynamic, 1) num_threads(P)")       \
        for (long c = 0; c < nchunks; ++c) {                                  \
            size_t lo = (size_t)c * per, hi = lo + per < n ? lo + per : n;    \
            size_t h[256];                                                    \
            memset(h, 0, sizeof h);                                           \
            for (size_t i = lo; i < hi; ++i) h[DG(&src[i], d)] += 1;          \
            memcpy(bins + (size_t)c * 256, h, sizeof h);                      \
        }                                                                     \
        size_t prefix = 0;                                                    \
        for (int g = 0; g < 256; ++g)                                         \
            for (long c = 0; c < nchunks; ++c) {                              \
                size_t next = prefix + bins[(size_t)c * 256 + g];             \
                bins[(size_t)c * 256 + g] = prefix;                           \
                prefix = next;                                                \
            }                                                                 \
        _Pragma("omp parallel for schedule(dynamic, 1) num_threads(P)")       \
        for (long c = 0; c < nchunks; ++c) {                                  \
            size_t lo = (size_t)c * per, hi = lo + per < n ? lo + per : n;    \
            size_t *starts = bins + (size_t)c * 256;                          \
            for (size_t i = lo; i < hi; ++i) {                                \
                unsigned g = DG(&src[i], d);                                  \
                memcpy(&dst[starts[g]], &src[i], sizeof(T));                  \
                starts[g] += 1;                                               \
            }                                                                 \
        }                                                                     \
    }                                                                         \
    if ((ND) % 2 == 1) memcpy(data, temp, n * sizeof(T));                     \
    free(bins); free(temp);                                                   \
    return 0;                                                                 \
}

#define DEF_ALL(NAME, T, ND, DG) DEF_SORT0(NAME, T, ND, DG) DEF_SORT_MT(NAME, T, ND, DG) DEF_SORT4_MT(NAME, T, ND, DG)

/* NUMBER_OF_DIGITS = byte width of the key (radix_digits.rs:8,16,24,32,40,...);
 * tuples take the key's (radix_digits.rs:131). */
DEF_ALL(u8, uint8_t, 1, dg_u8)
DEF_ALL(u16, uint16_t, 2, dg_u16)
DEF_ALL(u32, uint32_t, 4, dg_u32)
DEF_ALL(u64, uint64_t, 8, dg_u64)
DEF_ALL(u128, u128_t, 16, dg_u128)
DEF_ALL(i8, int8_t, 1, dg_i8)
DEF_ALL(i16, int16_t, 2, dg_i16)
DEF_ALL(i32, int32_t, 4, dg_i32)
DEF_ALL(i64, int64_t, 8, dg_i64)
DEF_ALL(i128, i128_t, 16, dg_i128)
DEF_ALL(f32, uint32_t, 4, dg_f32)
DEF_ALL(f64, uint64_t, 8, dg_f64)
DEF_ALL(p32, pair32_t, 4, dg_p32)
DEF_ALL(p64, pair64_t, 8, dg_p64)

#define DISPATCH(FN, ...)                                                     \
    switch (kind) {                                                           \
    case K_U8:   return FN##_u8((uint8_t *)data, __VA_ARGS__);                \
    case K_U16:  return FN##_u16((uint16_t *)data, __VA_ARGS__);              \
    case K_U32:  return FN##_u32((uint32_t *)data, __VA_ARGS__);              \
    case K_U64:  return FN##_u64((uint64_t *)data, __VA_ARGS__);              \
    case K_U128: return FN##_u128((u128_t *)data, __VA_ARGS__);               \
    case K_I8:   return FN##_i8((int8_t *)data, __VA_ARGS__);                 \
    case K_I16:  return FN##_i16((int16_t *)data, __VA_ARGS__);               \
    case K_I32:  return FN##_i32((int32_t *)data, __VA_ARGS__);               \
    case K_I64:  return FN##_i64((int64_t *)data, __VA_ARGS__);               \
    case K_I128: return FN##_i128((i128_t *)data, __VA_ARGS__);               \
    case K_F32:  return FN##_f32((uint32_t *)data, __VA_ARGS__);              \
    case K_F64:  return FN##_f64((uint64_t *)data, __VA_ARGS__);              \
    case K_PAIR_U32: return FN##_p32((pair32_t *)data, __VA_ARGS__);          \
    case K_PAIR_U64: return FN##_p64((pair64_t *)data, __VA_ARGS__);          \
    default: return -2;                                                       \
    }

/* Exported entry points (ctypes-loaded by oracle/cpu_oracle.py). */
int oracle_sort0(void *data, size_t n, int kind) { DISPATCH(sort0, n) }
int oracle_sort_mt(void *data, size_t n, int kind, int threads) { DISPATCH(sort_mt, n, threads) }
int oracle_sort4_mt(void *data, size_t n, int kind, int threads) { DISPATCH(sort4_mt, n, threads) }

/* get_digit for one element, for the known-answer tests. */
int oracle_get_digit(const void *e, int kind, unsigned index) {
    switch (kind) {
    case K_U8:   return (int)dg_u8((const uint8_t *)e, index);
    case K_U16:  return (int)dg_u16((const uint16_t *)e, index);
    case K_U32:  return (int)dg_u32((const uint32_t *)e, index);
    case K_U64:  return (int)dg_u64((const uint64_t *)e, index);
    case K_U128: return (int)dg_u128((const u128_t *)e, index);
    case K_I8:   return (int)dg_i8((const int8_t *)e, index);
    case K_I16:  return (int)dg_i16((const int16_t *)e, index);
    case K_I32:  return (int)dg_i32((const int32_t *)e, index);
    case K_I64:  return (int)dg_i64((const int64_t *)e, index);
    case K_I128: return (int)dg_i128((const i128_t *)e, index);
    case K_F32:  return (int)dg_f32((const uint32_t *)e, index);
    case K_F64:  return (int)dg_f64((const uint64_t *)e, index);
    default: return -2;
    }
}

/* ---- general (T, U) records: key = field .0 at byte offset key_off ------------------
 * radix_digits.rs:126-136 implements RadixDigits for ANY (T, U) with T: RadixDigits: the
 * digit is the key field's, the whole tuple is moved (mod.rs:133,145,158).  Sequential LSD,
 * the statement of mod.rs:182-212 with memcpy'd elements of rec_bytes bytes; `kind` names
 * the key type.  Returns 0, -2 for an unknown kind, -3 for a key that does not fit. */
static size_t kind_bytes(int kind) {
    switch (kind) {
    case K_U8: case K_I8: return 1;
    case K_U16: case K_I16: return 2;
    case K_U32: case K_I32: case K_F32: return 4;
    case K_U64: case K_I64: case K_F64: return 8;
    case K_U128: case K_I128: return 16;
    default: return 0;
    }
}

int oracle_sort0_records(void *data, size_t n, size_t rec_bytes, size_t key_off, int kind) {
    const size_t kb = kind_bytes(kind);
    if (kb == 0) return -2;
    if (key_off + kb > rec_bytes) return -3;
    if (n == 0) return 0;
    unsigned char *a = (unsigned char *)data;
    unsigned char *t = (unsigned char *)malloc(n * rec_bytes);
    if (!t) return -1;
    unsigned char *src = a, *dst = t;
    for (unsigned d = 0; d < (unsigned)kb; ++d) {           /* NUMBER_OF_DIGITS = size_of::<T>() */
        size_t hist[256] = {0};
        for (size_t i = 0; i < n; ++i) {                     /* mod.rs:191-194 */
            unsigned char key[16];
            memcpy(key, src + i * rec_bytes + key_off, kb);  /* the tuple may sit at any alignment */
            hist[oracle_get_digit(key, kind, d)]++;
        }
        size_t run = 0;                                      /* mod.rs:195-202 */
        for (int g = 0; g < 256; ++g) { size_t c = hist[g]; hist[g] = run; run += c; }
        for (size_t i = 0; i < n; ++i) {                     /* mod.rs:203-207 */
            unsigned char key[16];
            memcpy(key, src + i * rec_bytes + key_off, kb);
            memcpy(dst + hist[oracle_get_digit(key, kind, d)]++ * rec_bytes, src + i * rec_bytes, rec_bytes);
        }
        unsigned char *x = src; src = dst; dst = x;          /* mod.rs:186-190 */
    }
    if (src != a) memcpy(a, src, n * rec_bytes);             /* odd digit count, mod.rs:209-211 */
    free(t);
    return 0;
}

/* ---- input generator (bench.py's CPU arms) ----------------------------------------
 * The UNIFORM family of SURVEY.md 8(d): key i = low bytes of mix64(seed ^ mix64(i)),
 * mix64 = the splitmix64 finaliser; the same stream as radix_sort_b200.distr.uniform_host
 * and the device generator (inputs only: the reference draws from an unseeded thread_rng,
 * main.rs:27).  key_bytes in {1, 2, 4, 8}; 16: lo = u, hi = mix64(u). */
static inline uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

int oracle_generate_uniform(void *out, size_t n, int key_bytes, uint64_t seed, uint64_t index_base) {
    if (key_bytes != 1 && key_bytes != 2 && key_bytes != 4 && key_bytes != 8 && key_bytes != 16) return -2;
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; ++i) {
        const uint64_t u = mix64(seed ^ mix64(index_base + i));
        switch (key_bytes) {
        case 1: ((uint8_t *)out)[i] = (uint8_t)u; break;
        case 2: ((uint16_t *)out)[i] = (uint16_t)u; break;
        case 4: ((uint32_t *)out)[i] = (uint32_t)u; break;
        case 8: ((uint64_t *)out)[i] = u; break;
        default: ((uint64_t *)out)[2 * i] = u; ((uint64_t *)out)[2 * i + 1] = mix64(u); break;
        }
    }
    return 0;
}

/* available_parallelism() stand-in (mod.rs:66-70). */
int oracle_available_parallelism(void) {
#ifdef _OPENMP
    return omp_get_num_procs();
#else
    return 1;
#endif
}
