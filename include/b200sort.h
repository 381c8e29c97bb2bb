/*
 * b200sort.h -- C ABI of libb200sort.so, the B200 (sm_100a) LSD radix sort that
 * replaces the hot path of jgrodzki/radix_sort.
 *
 * The reference exposes the sort as a Rust trait method, not an FFI:
 *     pub trait RadixSort<T: RadixDigits> { fn radix_sort(&mut self); }
 *                                   (src/radix_sort/mod.rs:18-20, impl at :61-176)
 * keyed by `RadixDigits::get_digit`  (src/radix_sort/radix_digits.rs:1-136).
 * This header is the FFI layer a Rust `impl RadixSort<T> for [T]` binds to
 * (see INTEGRATION.md for the extern "C" block and the build.rs).
 *
 * Conventions: plain C, no C++ or torch types; every entry point returns an
 * int status (0 = ok) and never throws; the message for the calling thread's
 * last failure is b200sort_last_error().  There is NO CPU fallback: without a
 * CUDA device every compute entry point returns B200SORT_ERR_CUDA.
 * All entry points are re-entrant: concurrent calls from several host threads
 * on different buffers are safe (the reference's tests run that way under
 * libtest, src/radix_sort/tests.rs).
 */
#ifndef B200SORT_H
#define B200SORT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200SORT_VERSION 100 /* 0.1.0 */

/* Built-in key kinds = the RadixDigits impls of radix_digits.rs:7-124.
 * usize/isize (radix_digits.rs:47-53,95-101) are U64/I64 on 64-bit targets. */
typedef enum b200sort_kind {
    B200SORT_U8 = 0,   /* radix_digits.rs:7-13   */
    B200SORT_U16 = 1,  /* radix_digits.rs:15-21  */
    B200SORT_U32 = 2,  /* radix_digits.rs:23-29  */
    B200SORT_U64 = 3,  /* radix_digits.rs:31-37  */
    B200SORT_U128 = 4, /* radix_digits.rs:39-45  */
    B200SORT_I8 = 5,   /* radix_digits.rs:55-61  */
    B200SORT_I16 = 6,  /* radix_digits.rs:63-69  */
    B200SORT_I32 = 7,  /* radix_digits.rs:71-77  */
    B200SORT_I64 = 8,  /* radix_digits.rs:79-85  */
    B200SORT_I128 = 9, /* radix_digits.rs:87-93  */
    B200SORT_F32 = 10, /* radix_digits.rs:106-114 */
    B200SORT_F64 = 11, /* radix_digits.rs:116-124 */
    B200SORT_KIND_COUNT = 12
} b200sort_kind;

typedef enum b200sort_status {
    B200SORT_OK = 0,
    B200SORT_ERR_INVALID_ARG = 1,      /* null pointer, misaligned buffer, short workspace */
    B200SORT_ERR_UNSUPPORTED_KIND = 2, /* a user-defined RadixDigits impl cannot run on a GPU */
    B200SORT_ERR_CUDA = 3,             /* any CUDA runtime failure, incl. "no device" */
    B200SORT_ERR_OOM = 4               /* device or pinned-host allocation failed */
} b200sort_status;

/* flags for the *_ex entry points */
#define B200SORT_FLAG_NO_SKIP 1u /* run every digit pass even when a digit is constant */

int b200sort_version(void);

/* Message of the calling thread's last failing call ("" if none). */
const char *b200sort_last_error(void);

/* NUMBER_OF_DIGITS / size in bytes of a key kind (radix_digits.rs:8,16,...,117);
 * 0 for an unknown kind. */
size_t b200sort_key_bytes(int kind);

/* Bytes of device workspace b200sort_keys_device needs for n keys of `kind`:
 * the ping-pong buffer the reference allocates at mod.rs:71-83 (n * key bytes)
 * plus histograms, bucket bases and look-back state. 0 for an unknown kind. */
size_t b200sort_temp_bytes(size_t n, int kind);

/* Device-resident sort: replaces the body of RadixSort::radix_sort
 * (mod.rs:61-176) for keys already in HBM.  Sorts d_keys[0..n) in place
 * (ascending in the order get_digit defines: Ord for integers, total_cmp for
 * floats), asynchronously on `stream` (a cudaStream_t; NULL = default stream).
 * d_tmp/tmp_bytes: workspace of at least b200sort_temp_bytes(n, kind) bytes,
 * 256-byte aligned; contents are clobbered.  No host synchronisation, no
 * allocation: safe to capture into a CUDA graph.  n == 0 is a no-op (the
 * reference panics there, mod.rs:92; documented deviation). */
int b200sort_keys_device(void *d_keys, void *d_tmp, size_t tmp_bytes, size_t n,
                         int kind, void *stream);
int b200sort_keys_device_ex(void *d_keys, void *d_tmp, size_t tmp_bytes, size_t n,
                            int kind, void *stream, unsigned flags);

/* The same device-resident sort for callers that have no way to provide d_tmp.
 * (No counterpart in the reference: it sorts host slices and takes its temp buffer from
 * the Rust allocator, mod.rs:71-83; this exists only because of how GPU hosts are built.)
 *
 * Who that is: a host that binds nothing but this library and is handed device pointers
 * by someone else (the Rust shim links libb200sort.so only -- it has no cudaMalloc to
 * call); a plugin living inside a process whose own allocator has already taken nearly
 * all of HBM, where a cudaMalloc from a guest library fails or steals from the owner's
 * budget; and loops that cannot tolerate the device-wide synchronisation that
 * cudaMalloc / cudaFree imply.  For them the workspace is a fixed array inside the
 * library image: B200SORT_NOALLOC_SCRATCH_BYTES of HBM on every GPU that loads the
 * library, held until the process exits, whether or not this entry point is ever called.
 *
 * Sorts d_keys[0..n) in place like b200sort_keys_device, asynchronously on `stream`, on
 * the device that is current in the calling thread.  The call itself never allocates or
 * frees device memory, neither directly nor through a library routine.
 * Limit: b200sort_temp_bytes(n, kind) <= B200SORT_NOALLOC_SCRATCH_BYTES (about 1.0 M
 * 16-byte, 2.0 M 8-byte, 3.9 M 4-byte, 1.5 M 2-byte keys); anything larger is refused with
 * B200SORT_ERR_INVALID_ARG and a message naming both numbers -- nothing is launched.
 * The first call on each device looks up the array's device address
 * (cudaGetSymbolAddress, once per device, cached); make that first call outside any
 * stream capture.
 * The array is shared by every call on that device and keeps whatever the previous sort
 * left in it; each sort re-zeroes the counters it depends on, so back-to-back calls on
 * one stream are fine.  Two calls whose kernels can overlap on the GPU (different
 * streams without an event between them, or different host threads) race on the array
 * and give garbage: order them yourself. */
#define B200SORT_NOALLOC_SCRATCH_BYTES (16u << 20) /* 16 MiB */
int b200sort_keys_device_noalloc(void *d_keys, size_t n, int kind, void *stream);

/* The device-resident sort with the temp buffer taken and returned INSIDE the call, as the
 * reference does (mod.rs:71-83 allocates, :176 drops; its timer, main.rs:32-34, covers both).
 * The workspace comes from a stream-ordered memory pool the library owns (one per device,
 * cudaMallocFromPoolAsync / cudaFreeAsync on `stream`, freed blocks kept for the next call), so
 * only the first call of a size pays a real allocation: cudaMalloc + cudaFree around
 * b200sort_keys_device cost 8.7 ms per 1 B-key sort (24 %).  Asynchronous on `stream`, on the
 * current device; B200SORT_ERR_OOM when the pool cannot grow.  The pool's idle memory is
 * returned to the driver by b200sort_release_cached_memory(). */
int b200sort_keys_device_pooled(void *d_keys, size_t n, int kind, void *stream);

/* Drop-in for `slice.radix_sort()` on a host slice (call sites
 * tests.rs:30,...,166; main.rs:33): copies keys to the current device, sorts,
 * copies back into `keys`, blocks until done.  Owns its device memory. */
int b200sort_keys_host(void *keys, size_t n, int kind);

/* Host slices of at least this many bytes (default 64 MiB) take the pipelined
 * form of b200sort_keys_host: the keys are split into 16 key ranges on the
 * device (K7, splitters sampled from the host slice), the ranges are sorted one
 * after the other and each is copied back while the next is being sorted.  Same
 * result, the device->host copy overlaps the sort.  0 = always, SIZE_MAX = never
 * (tests use both). */
int b200sort_set_host_pipeline_min_bytes(size_t bytes);

/* The host entry points keep the device blocks they staged through (2 x data +
 * workspace: ~17-25 GB after one 1 B-key u64 call) for the next call, and the
 * communicator keeps its receive buffers.  This frees every block of the host-path
 * pool that is not in use right now (all devices); the next call allocates again.
 * Returns B200SORT_OK. */
int b200sort_release_cached_memory(void);

/* Key-value records, key = field 0 of a `(K, V)` tuple (radix_digits.rs:126-136;
 * call sites tests.rs:182, main.rs:112,123).  Stable.  rec_bytes is the whole
 * record (8 for (u32,u32), 16 for (u64,u64)); the key of `kind` sits at byte
 * offset key_off inside each record. */
size_t b200sort_records_temp_bytes(size_t n, size_t rec_bytes, int kind);
int b200sort_records_device(void *d_recs, void *d_tmp, size_t tmp_bytes, size_t n,
                            size_t rec_bytes, size_t key_off, int kind, void *stream);
int b200sort_records_host(void *recs, size_t n, size_t rec_bytes, size_t key_off, int kind);

/* ---- instrumentation (bench / tests; not on the reference's API) ---------- */

/* Number of kernels the library has launched from this process so far
 * (memset nodes not counted).  bench.py reports the delta as gpu_launches. */
unsigned long long b200sort_launch_count(void);

/* Test hooks for the K3 slow path.  A tile whose order check fails is re-ranked
 * with ballots and rewritten; that has never been observed on B200, so tests
 * force it: b200sort_debug_force_rerank(1) makes every tile of every later sort
 * of this process take the slow path (0 restores).  b200sort_debug_reranked_tiles
 * synchronises `stream` and returns in *out how many tiles of the LAST sort run
 * on d_tmp (same n, kind; rec_bytes = 0 for key sorts, 8 / 16 for record sorts)
 * took it. */
void b200sort_debug_force_rerank(int on);
/* Slices of at most 4096 elements are sorted by one launch of one CTA (all digit
 * passes in shared memory) instead of the K1-K2-K3 sequence; 0 turns that off so
 * that tests can drive the big path at small sizes too (1 restores). */
void b200sort_debug_small_sort(int on);
/* Key-only sorts of 1- and 2-byte keys (u8, i8, u16, i16; counting_sort, mod.rs:40-59) need no
 * ranking: equal keys are bit-identical, so the library takes one value histogram and writes
 * every value count[v] times (csrc/narrow_kernels.cuh).  0 sends these widths through the
 * onesweep passes instead (as B200SORT_FLAG_NO_SKIP and the timed entry point always do), so that
 * tests reach both; 1 restores. */
void b200sort_debug_narrow_fill(int on);
int b200sort_debug_reranked_tiles(void *d_tmp, size_t tmp_bytes, size_t n, int kind, size_t rec_bytes, void *stream,
                                  unsigned *out);

/* Kernel-tuning override for sweeps: K3 variant index per element width.
 * key_bytes = 1, 2, 4, 8, 16 selects the key-only table of that width, 108 / 116
 * the tables of 8- / 16-byte records.  -1 restores the default.  Returns
 * B200SORT_ERR_INVALID_ARG if unknown.  b200sort_num_variants: size of a table. */
int b200sort_set_variant(int key_bytes, int variant);
int b200sort_num_variants(int key_bytes);

/* Per-phase timing of one device sort, measured with CUDA events on `stream`
 * (synchronises!).  out_ms[0] = K1 histogram, out_ms[1] = K2 scan,
 * out_ms[2 + p] = digit pass p (0 if skipped), out_ms[2 + D] = copy-back,
 * out_ms[3 + D] = total.  out_len must be >= D + 4. */
int b200sort_keys_device_timed(void *d_keys, void *d_tmp, size_t tmp_bytes, size_t n,
                               int kind, void *stream, unsigned flags,
                               float *out_ms, int out_len);

/* ---- synthetic inputs and on-device verification (K5 / K6) ---------------- */

/* distribution families of src/distr.rs, restated as counter-based generators
 * (value = f(seed, index)), identical on CPU (radix_sort_b200/distr.py). */
typedef enum b200sort_distr {
    B200SORT_DISTR_UNIFORM = 0,    /* Standard / KeyUniform     distr.rs:40-52  */
    B200SORT_DISTR_GEOMETRIC = 1,  /* MyExp(lambda)=Geometric   distr.rs:4-38   */
    B200SORT_DISTR_ZIPF = 2,       /* ZipfU32/ZipfU64(s)        distr.rs:54-76,108-130 */
    B200SORT_DISTR_STEP = 3,       /* StepUniformU32/U64(m)     distr.rs:78-106,132-160 */
    B200SORT_DISTR_ASCENDING = 4,  /* presorted (builder's; not in distr.rs) */
    B200SORT_DISTR_DESCENDING = 5, /* presorted, reversed (builder's) */
    B200SORT_DISTR_FLOAT_TEST = 6  /* Uniform[-1,1) + specials  tests.rs:135-143 */
} b200sort_distr;

/* Fill d_keys[0..n) with keys index_base .. index_base+n of the stream
 * (seed, distr, param).  param: lambda (GEOMETRIC), s (ZIPF), m (STEP). */
int b200sort_generate(void *d_keys, size_t n, int kind, int distr, double param,
                      uint64_t seed, uint64_t index_base, void *stream);

/* Order-independent multiset fingerprint of d_keys[0..n):
 * out[0] = n, out[1] = sum mix64(key) mod 2^64, out[2] = xor mix64(key),
 * out[3] = sum key-low-64 mod 2^64.  d_out: 4 x uint64 on the device. */
int b200sort_fingerprint(const void *d_keys, size_t n, int kind, uint64_t *d_out,
                         void *stream);

/* Count adjacent pairs that are out of order under get_digit's order:
 * *d_out (uint64 on the device) = #{i : key[i] > key[i+1]}. 0 <=> sorted. */
int b200sort_count_unsorted(const void *d_keys, size_t n, int kind, uint64_t *d_out,
                            void *stream);

/* ---- multi-GPU building blocks (one process per GPU: radix_sort_b200/mgpu.py drives
 * them; the exchange is K7 storing into the destinations' receive buffers over NVLink
 * -- b200sort_partition_exchange_device below -- with K7 into a send buffer + the
 * caller's NCCL all-to-all as the fallback when peer memory cannot be mapped) ---- */

/* K7: partition of d_in[0..n) by destination rank (the order inside a destination is
 * NOT specified: every destination is sorted next, and equal keys are bit-identical).  A key goes to rank
 * r = #{j : splitters[j] <= key}, 0 <= r <= nsplit (nsplit <= 15), compared in
 * get_digit's order,
 * where `splitters` are nsplit ascending keys of `kind` in HOST memory, given
 * as raw key bit patterns.  Writes the keys grouped by destination to d_out
 * and the per-destination counts (nsplit + 1 uint64) to d_counts.
 * Workspace: b200sort_temp_bytes(n, kind). */
int b200sort_partition_device(const void *d_in, void *d_out, void *d_tmp, size_t tmp_bytes,
                              size_t n, int kind, const void *splitters, int nsplit,
                              uint64_t *d_counts, void *stream);

/* K7 fused with the exchange: the same partition, split in two calls around the
 * caller's count exchange, storing every bucket straight into the receive
 * buffer of its destination GPU over NVLink (peer memory mapped into this
 * process: CUDA IPC / symmetric memory; P2P access must be enabled).
 *
 *   1. b200sort_partition_count_device: reads the keys once, writes the
 *      per-destination counts (nsplit + 1 uint64) to d_counts and leaves its
 *      state in d_tmp.
 *   2. the caller exchanges the counts (all_gather) and computes, for every
 *      destination r, the element offset in r's receive buffer at which this
 *      rank's keys start (= keys sent to r by lower ranks).
 *   3. b200sort_partition_exchange_device: SAME d_in, d_tmp, n, kind and
 *      splitters, d_tmp untouched in between.  peer_bases: nsplit + 1 device
 *      pointers in HOST memory, peer_bases[r] = rank r's receive buffer as
 *      addressable from this device (own rank: the local buffer).
 *      d_peer_offsets: nsplit + 1 uint64 element offsets in DEVICE memory.
 *      Returns when the stores are enqueued; the caller synchronises the
 *      ranks (barrier after stream completion) before reading receive buffers.
 * There is no reference counterpart (the reference is single-process); this
 * replaces the all-to-all of the unfused path (radix_sort_b200/mgpu.py). */
int b200sort_partition_count_device(const void *d_in, void *d_tmp, size_t tmp_bytes, size_t n, int kind,
                                    const void *splitters, int nsplit, uint64_t *d_counts, void *stream);
int b200sort_partition_exchange_device(const void *d_in, void *d_tmp, size_t tmp_bytes, size_t n, int kind,
                                       const void *splitters, int nsplit, void *const *peer_bases,
                                       const uint64_t *d_peer_offsets, void *stream);
/* K7 with no count phase at all: one read of d_in, every tile CLAIMS room for its keys of
 * bucket r with one system-scope atomicAdd on the u64 cursor at device address cursor_addrs[r]
 * (on this GPU or, peer-mapped, on bucket r's owner) and stores them at region_bases[r] + the
 * claimed element offset.  The order of the keys inside a region is unspecified (the region is
 * sorted next; equal keys are bit-identical).  region_caps[r] = elements the region holds: a
 * claim that would pass it is not stored and sets bit r of *d_flags; the cursor counts it all
 * the same, so after the call (and a barrier over all writers) a cursor holds the true size of
 * its bucket.  Cursors and *d_flags must be zero before the first writer starts.  No counts
 * travel, no all_gather, no second read of the shard.  splitters / region_bases /
 * cursor_addrs / region_caps: host arrays (nsplit, nsplit + 1, ...).  Workspace:
 * b200sort_temp_bytes(n, kind). */
int b200sort_partition_claim_device(const void *d_in, void *d_tmp, size_t tmp_bytes, size_t n, int kind,
                                    const void *splitters, int nsplit, void *const *region_bases,
                                    const uint64_t *cursor_addrs, const uint64_t *region_caps,
                                    uint32_t *d_flags, void *stream);

/* Step 3 in pieces, so that the receiver can start sorting the buckets of one
 * piece while the next piece is still crossing NVLink: stores only the buckets
 * whose bit is set in bucket_mask (bit r = bucket r); the other keys are read
 * and ranked but not written.  again = 0 for the first piece after the count
 * call, 1 for every later piece over the same d_tmp (re-arms the tile tickets
 * and look-back state the previous piece consumed).  Pieces over one d_tmp must
 * not overlap in time (enqueue them in stream order). */
int b200sort_partition_exchange_part_device(const void *d_in, void *d_tmp, size_t tmp_bytes, size_t n, int kind,
                                            const void *splitters, int nsplit, void *const *peer_bases,
                                            const uint64_t *d_peer_offsets, unsigned bucket_mask, int again,
                                            void *stream);

/* ---- all GPUs of one process behind one call --------------------------------
 * The reference is a single process (`slice.radix_sort()`, mod.rs:18-20,61); a
 * caller that holds shards on several GPUs of that process sorts them with one
 * communicator, no launcher and no NCCL: sampled splitters, K7 count, K7
 * scatter storing every key straight into its destination GPU's receive buffer
 * over NVLink (peer access is enabled by create), one local sort per GPU.
 *
 * b200sort_mgpu_create: ndev in [1, 16]; devs = CUDA device ordinals (NULL =
 *   0 .. ndev-1).  A device may be named more than once (several shards on one
 *   GPU).  Fails with B200SORT_ERR_CUDA when two devices have no peer access.
 * b200sort_mgpu_keys_device: d_shards[r] = n_in[r] keys of `kind` in device
 *   memory of devs[r]; they are read, not written.  On return d_out[r] points
 *   to n_out[r] sorted keys on devs[r]: the r-th contiguous range of the global
 *   order, so that d_out[0] ++ d_out[1] ++ ... is what the reference's sort
 *   gives for the concatenated input.  The buffers belong to the communicator
 *   and stay valid until its next sort or its destruction.  Blocking; the
 *   current device of the calling thread is restored.  One sort at a time per
 *   communicator (calls are serialised). */
typedef struct b200sort_comm b200sort_comm;
int b200sort_mgpu_create(b200sort_comm **comm, int ndev, const int *devs);
int b200sort_mgpu_keys_device(b200sort_comm *comm, const void *const *d_shards, const size_t *n_in, int kind,
                              void **d_out, size_t *n_out);
int b200sort_mgpu_destroy(b200sort_comm *comm);
/* The drop-in call on all devices of the communicator: `keys` is a HOST slice
 * of n keys (what `slice.radix_sort()` is called on, mod.rs:18-20), sorted in
 * place.  Device r copies the r-th n/ndev keys in over its own PCIe link, the
 * parts are sorted as one key set (b200sort_mgpu_keys_device), and device r
 * copies its range of the result back to its place in the slice.  That is the
 * route of a PINNED slice (cudaHostRegister / cudaHostAlloc); a pageable one,
 * slices below 32 MiB (b200sort_mgpu_set_host_min_bytes) and one-device
 * communicators go to b200sort_keys_host on the first device.  Blocking. */
int b200sort_mgpu_keys_host(b200sort_comm *comm, void *keys, size_t n, int kind);
int b200sort_mgpu_set_host_min_bytes(size_t bytes);

#ifdef __cplusplus
}
#endif
#endif /* B200SORT_H */
