// b200sort_api.cu -- the C ABI of libb200sort.so (include/b200sort.h): dispatch
// over key kinds, workspace layout, launch sequence of K1..K3, host-slice path.
//
// Replaces RadixSort::radix_sort for [T] (reference src/radix_sort/mod.rs:61-176).
// There is deliberately no CPU path in this file: every compute entry point
// either launches the sm_100a kernels or returns an error.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "sort_impl.cuh"
#include "host_staging.cuh"

using namespace b200;

namespace {
thread_local std::string g_last_error;
std::atomic<unsigned long long> g_launches{0};
std::atomic<int> g_variant[3][17];  // [family][element bytes]: selected variant + 1; 0 = default
}  // namespace

namespace b200 {
int fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}
int cuda_fail(cudaError_t e, const char* what) {
    return fail(e == cudaErrorMemoryAllocation ? B200SORT_ERR_OOM : B200SORT_ERR_CUDA,
                std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")");
}
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
std::atomic<int> g_force_rerank{0};
bool force_rerank() { return g_force_rerank.load(std::memory_order_relaxed) != 0; }
std::atomic<int> g_small_sort{1};
bool small_sort_enabled() { return g_small_sort.load(std::memory_order_relaxed) != 0; }
std::atomic<int> g_narrow_fill{1};
bool narrow_fill_enabled() { return g_narrow_fill.load(std::memory_order_relaxed) != 0; }
int variant_override(int family, int elem_bytes) {
    return g_variant[family][elem_bytes].load(std::memory_order_relaxed) - 1;
}
void set_last_error(const std::string& msg) { g_last_error = msg; }  // used by aux_kernels.cu

// instantiated by inst_front.cu, one translation unit per (family, width)
#define B200_DECL_FRONT(FAM, W)                                   \
    extern template int sort_front<FAM, W>(const SortCall&);      \
    extern template size_t temp_bytes_front<FAM, W>(size_t);
template <int FAM, int W> int sort_front(const SortCall&);
template <int FAM, int W> size_t temp_bytes_front(size_t);
B200_DECL_FRONT(FAM_KEYS, 1) B200_DECL_FRONT(FAM_KEYS, 2) B200_DECL_FRONT(FAM_KEYS, 4)
B200_DECL_FRONT(FAM_KEYS, 8) B200_DECL_FRONT(FAM_KEYS, 16)
B200_DECL_FRONT(FAM_RECORDS, 8) B200_DECL_FRONT(FAM_RECORDS, 16)
B200_DECL_FRONT(FAM_PARTITION, 1) B200_DECL_FRONT(FAM_PARTITION, 2) B200_DECL_FRONT(FAM_PARTITION, 4)
B200_DECL_FRONT(FAM_PARTITION, 8) B200_DECL_FRONT(FAM_PARTITION, 16)
}  // namespace b200

#define CU(call) B200_CU(call)

namespace {

// key kind -> transform of RadixDigits::get_digit (radix_digits.rs:7-124)
int kind_xf(int kind) {
    switch (kind) {
    case B200SORT_I8: case B200SORT_I16: case B200SORT_I32: case B200SORT_I64: case B200SORT_I128: return XF_SIGNED;
    case B200SORT_F32: case B200SORT_F64: return XF_FLOAT;
    default: return XF_NONE;
    }
}

template <int FAM>
int front_by_width(size_t w, const SortCall& c) {
    switch (w) {
    case 1: if constexpr (FAM != FAM_RECORDS) return sort_front<FAM, 1>(c); break;
    case 2: if constexpr (FAM != FAM_RECORDS) return sort_front<FAM, 2>(c); break;
    case 4: if constexpr (FAM != FAM_RECORDS) return sort_front<FAM, 4>(c); break;
    case 8: return sort_front<FAM, 8>(c);
    case 16: return sort_front<FAM, 16>(c);
    default: break;
    }
    return fail(B200SORT_ERR_UNSUPPORTED_KIND, "unsupported element width " + std::to_string(w));
}

template <int FAM>
size_t temp_by_width(size_t w, size_t n) {
    switch (w) {
    case 1: if constexpr (FAM != FAM_RECORDS) return temp_bytes_front<FAM, 1>(n); break;
    case 2: if constexpr (FAM != FAM_RECORDS) return temp_bytes_front<FAM, 2>(n); break;
    case 4: if constexpr (FAM != FAM_RECORDS) return temp_bytes_front<FAM, 4>(n); break;
    case 8: return temp_bytes_front<FAM, 8>(n);
    case 16: return temp_bytes_front<FAM, 16>(n);
    default: break;
    }
    return 0;
}

int keys_device_dispatch(void* d_keys, void* d_tmp, size_t tmp_bytes, size_t n, int kind,
                         cudaStream_t st, unsigned flags, float* out_ms, int out_len) {
    const size_t kb = b200sort_key_bytes(kind);
    if (kb == 0) return fail(B200SORT_ERR_UNSUPPORTED_KIND, "unknown key kind " + std::to_string(kind));
    SortCall c;
    c.elems = d_keys; c.ws = d_tmp; c.ws_bytes = tmp_bytes; c.n = n; c.st = st; c.flags = flags;
    c.xf = kind_xf(kind); c.out_ms = out_ms; c.out_len = out_len;
    return front_by_width<FAM_KEYS>(kb, c);
}

// ---- records: (K, V) tuples, key = one field (radix_digits.rs:126-136) ------------
int check_record_shape(size_t rec_bytes, size_t key_off, int kind) {
    const size_t kb = b200sort_key_bytes(kind);
    if (kb == 0) return fail(B200SORT_ERR_UNSUPPORTED_KIND, "unknown key kind " + std::to_string(kind));
    const bool ok = (rec_bytes == 8 && kb == 4 && (key_off == 0 || key_off == 4)) ||
                    (rec_bytes == 16 && kb == 8 && (key_off == 0 || key_off == 8));
    if (!ok)
        return fail(B200SORT_ERR_UNSUPPORTED_KIND,
                    "record shape not supported: only (K,V) with 4-byte key in an 8-byte record or 8-byte key in a "
                    "16-byte record");
    return B200SORT_OK;
}

// ---- device memory pool for the host-slice entry points ---------------------------
// Keeps freed blocks for reuse so that repeated host sorts (bench loops, a test
// suite) do not pay cudaMalloc/cudaFree every call.  Mutex-guarded: concurrent
// callers each get their own block.
struct PoolBlock { void* p; size_t bytes; int dev; bool busy; };
std::mutex g_pool_mu;
std::vector<PoolBlock> g_pool;

int pool_acquire(size_t bytes, void** out) {
    int dev = 0;
    CU(cudaGetDevice(&dev));
    {
        std::lock_guard<std::mutex> lk(g_pool_mu);
        int best = -1;
        for (int i = 0; i < (int)g_pool.size(); ++i)
            if (!g_pool[i].busy && g_pool[i].dev == dev && g_pool[i].bytes >= bytes &&
                (best < 0 || g_pool[i].bytes < g_pool[best].bytes))
                best = i;
        if (best >= 0) {
            g_pool[best].busy = true;
            *out = g_pool[best].p;
            return B200SORT_OK;
        }
        // no fit: drop idle blocks of this device (they are too small) before growing
        for (int i = (int)g_pool.size() - 1; i >= 0; --i)
            if (!g_pool[i].busy && g_pool[i].dev == dev) {
                cudaFree(g_pool[i].p);
                g_pool.erase(g_pool.begin() + i);
            }
    }
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc(device staging + workspace)");
    std::lock_guard<std::mutex> lk(g_pool_mu);
    g_pool.push_back({p, bytes, dev, true});
    *out = p;
    return B200SORT_OK;
}
void pool_release(void* p) {
    std::lock_guard<std::mutex> lk(g_pool_mu);
    for (auto& b : g_pool)
        if (b.p == p) b.busy = false;
}

struct StreamGuard {
    cudaStream_t st = nullptr;
    ~StreamGuard() { if (st) cudaStreamDestroy(st); }
};

// The plain round trip runs on one non-blocking stream per (host thread, device), created
// on first use and kept: cudaStreamCreate + cudaStreamDestroy per call was a third of the
// 143 us a 1000-key host slice took.  A thread runs one host sort at a time, so the stream
// is never shared; it is deliberately not destroyed at thread exit (the driver may be gone).
cudaError_t thread_stream(cudaStream_t* out) {
    thread_local cudaStream_t streams[32] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    cudaStream_t& s = streams[dev & 31];
    if (!s) {
        e = cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
        if (e != cudaSuccess) { s = nullptr; return e; }
    }
    *out = s;
    return cudaSuccess;
}

// copy in, sort, copy out; `sort` is called with (device elems, workspace, workspace bytes, stream)
template <typename SortFn>
int host_roundtrip(void* host, size_t bytes, size_t ws_bytes, SortFn sort) {
    const size_t data_bytes = align_up(bytes, 256);
    void* block = nullptr;
    int rc = pool_acquire(data_bytes + ws_bytes, &block);
    if (rc) return rc;
    struct { cudaStream_t st = nullptr; } sg;
    cudaError_t e = thread_stream(&sg.st);
    if (e != cudaSuccess) { pool_release(block); return cuda_fail(e, "cudaStreamCreate"); }
    unsigned char* d = static_cast<unsigned char*>(block);
    // large pageable slices (a plain Vec<T>) go through the library's pinned bounce
    // buffers and copy threads (host_staging.cuh); everything else is one DMA each way
    Bounce bounce;
    // (B200SORT_STAGE_MIN_BYTES: sweep knob for where the bounce path starts to beat the driver's own pageable copies)
    static const size_t stage_min = [] {
        const char* e = std::getenv("B200SORT_STAGE_MIN_BYTES");
        // 16 MiB: a 40 MB pageable u32 slice takes 4.6 ms staged against 6.9 ms through the driver's pageable copies; at
        // 4 MiB an 8 MB slice gets slower (2.2 against 1.35 ms) -- profiles/r03_host_latency_stage_min.jsonl
        return e ? (size_t)std::strtoull(e, nullptr, 10) : (size_t)16 << 20;
    }();
    const bool staged = bytes >= stage_min && is_pageable(host);
    if (staged) e = bounce_acquire(&bounce);
    if (e != cudaSuccess) { pool_release(block); return cuda_fail(e, "cudaHostAlloc(bounce buffers)"); }
    if (staged) {
        size_t issued = 0;
        e = staged_copy_in(bounce, d, static_cast<const unsigned char*>(host), bytes, sg.st, &issued);
    } else {
        e = cudaMemcpyAsync(d, host, bytes, cudaMemcpyHostToDevice, sg.st);
    }
    if (e == cudaSuccess) {
        rc = sort(d, d + data_bytes, ws_bytes, sg.st);
        if (rc == B200SORT_OK) {
            if (staged) {
                StagedCopyOut out(bounce, sg.st);
                e = out.add(static_cast<unsigned char*>(host), d, bytes);
                if (e == cudaSuccess) e = out.finish();
            } else {
                e = cudaMemcpyAsync(host, d, bytes, cudaMemcpyDeviceToHost, sg.st);
            }
        }
    }
    const cudaError_t e2 = cudaStreamSynchronize(sg.st);
    if (staged) bounce_release(bounce);
    pool_release(block);
    if (rc) return rc;
    if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpyAsync");
    if (e2 != cudaSuccess) return cuda_fail(e2, "cudaStreamSynchronize");
    return B200SORT_OK;
}


// ---- pipelined host path (keys, large inputs) ---------------------------------------
// The plain round trip serialises copy-in, sort and copy-out: 8 GB + 8 GB over
// PCIe around a 39 ms sort.  The copy-in cannot overlap the sort (every pass
// needs all keys), but the copy-out can: split the keys into 16 key ranges
// (splitters sampled from the host buffer), sort the ranges one after the other,
// and start the device->host copy of a range as soon as it is sorted, on a second
// stream.  The split itself hides under the copy-in: the keys arrive in 16 chunks
// and each chunk is split by one K7 run as it lands, into its own slice of a
// second buffer.  After the last byte only that chunk's split (0.6 ms at 1 B u64)
// is left; a range is then 16 pieces, gathered into the freed input buffer by
// device copies right before its sort (0.3 ms per range, behind the copy-out of
// the range before it).  The gathered ranges are contiguous and in order, so the
// host buffer ends up exactly as the plain path leaves it.
std::atomic<size_t> g_pipeline_min_bytes{(size_t)64 << 20};
constexpr int PIPE_GROUPS = MAX_SPLITTERS + 1;

template <typename U>
void host_splitters(const void* host, size_t n, int xf, U* out_raw) {
    const unsigned char* base = static_cast<const unsigned char*>(host);
    const size_t S = std::min<size_t>(n, (size_t)1 << 14);
    const size_t stride = n / S;
    std::vector<U> s(S);
    for (size_t i = 0; i < S; ++i) {
        U raw;
        std::memcpy(&raw, base + (i * stride + (size_t)(mix64(i) % stride)) * sizeof(U), sizeof(U));
        s[i] = to_ordered_rt<U>(raw, xf);
    }
    std::sort(s.begin(), s.end(), [](const U& a, const U& b) { return ordered_less<U>(a, b); });
    for (int j = 0; j < MAX_SPLITTERS; ++j) out_raw[j] = from_ordered_rt<U>(s[S * (size_t)(j + 1) / PIPE_GROUPS], xf);
}

struct Trace {  // B200SORT_TRACE=1: wall-clock marks of the host path on stderr
    bool on = std::getenv("B200SORT_TRACE") != nullptr;
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    void mark(const char* what) {
        if (!on) return;
        const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        std::fprintf(stderr, "[b200sort] %8.2f ms  %s\n", ms, what);
    }
};

int keys_host_pipelined(void* host, size_t n, int kind) {
    Trace tr;
    const size_t kb = b200sort_key_bytes(kind);
    const size_t bytes = n * kb;
    const size_t data = align_up(bytes, 256);
    constexpr int CHUNKS = 16;
    const size_t per_chunk = ((n + CHUNKS - 1) / CHUNKS + 4095) / 4096 * 4096;  // whole 16-byte vectors for K1
    const size_t ws_part = align_up(temp_by_width<FAM_PARTITION>(kb, per_chunk), 256);
    const size_t counts_bytes = align_up((size_t)CHUNKS * PIPE_GROUPS * 8, 256);
    // workspace of the range sorts, sized for ranges up to 1.25x the even share; a more
    // skewed split takes a second block once the counts are known
    const size_t ws_sort_cap = align_up(b200sort_temp_bytes(n / PIPE_GROUPS + n / (4 * PIPE_GROUPS) + 4096, kind), 256);
    void* block = nullptr;
    int rc = pool_acquire(2 * data + ws_part + counts_bytes + ws_sort_cap, &block);
    if (rc) return rc;
    struct Cleanup {
        void* block; void* extra = nullptr; cudaStream_t s = nullptr, c = nullptr; cudaEvent_t ev[PIPE_GROUPS] = {};
        Bounce bounce; bool have_bounce = false;
        ~Cleanup() {
            if (s) cudaStreamSynchronize(s);
            if (c) cudaStreamSynchronize(c);
            for (auto e : ev) if (e) cudaEventDestroy(e);
            if (s) cudaStreamDestroy(s);
            if (c) cudaStreamDestroy(c);
            if (have_bounce) bounce_release(bounce);
            if (extra) pool_release(extra);
            pool_release(block);
        }
    } cl{block};
    // a pageable slice (a plain Vec<T>) goes through pinned bounce buffers filled and
    // emptied by the library's copy threads (host_staging.cuh); a pinned one is DMA'd directly
    const bool staged = is_pageable(host);
    if (staged) {
        CU(bounce_acquire(&cl.bounce));
        cl.have_bounce = true;
    }
    CU(cudaStreamCreateWithFlags(&cl.s, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&cl.c, cudaStreamNonBlocking));
    unsigned char* d_in = static_cast<unsigned char*>(block);
    unsigned char* d_out = d_in + data;
    unsigned char* d_ws = d_out + data;
    uint64_t* d_counts = reinterpret_cast<uint64_t*>(d_ws + ws_part);
    unsigned char* ws_sort = d_ws + ws_part + counts_bytes;
    size_t ws_sort_bytes = ws_sort_cap;

    tr.mark("streams + device block ready");
    // the copy-in goes in chunks on the copy stream; while the DMA runs the host picks
    // the splitters (1-2 ms), then every chunk is split by destination range (one K7 run
    // into its own slice of d_out) on the compute stream as soon as it has landed: when
    // the last byte arrives only the split of the last chunk is left to do
    cudaEvent_t landed[CHUNKS] = {};
    struct EvGuard { cudaEvent_t* e; ~EvGuard() { for (int i = 0; i < CHUNKS; ++i) if (e[i]) cudaEventDestroy(e[i]); } } eg{landed};
    if (!staged) {
        for (int c = 0; c < CHUNKS; ++c) {
            const size_t off = (size_t)c * per_chunk;
            if (off >= n) break;
            const size_t cnt = std::min(per_chunk, n - off);
            CU(cudaMemcpyAsync(d_in + off * kb, static_cast<const unsigned char*>(host) + off * kb, cnt * kb, cudaMemcpyHostToDevice, cl.c));
            CU(cudaEventCreateWithFlags(&landed[c], cudaEventDisableTiming));
            CU(cudaEventRecord(landed[c], cl.c));
        }
        tr.mark("copy-in enqueued");
    }
    alignas(16) unsigned char spl[MAX_SPLITTERS * 16];
    const int xf = kind_xf(kind);
    switch (kb) {
    case 1: host_splitters<unsigned char>(host, n, xf, reinterpret_cast<unsigned char*>(spl)); break;
    case 2: host_splitters<unsigned short>(host, n, xf, reinterpret_cast<unsigned short*>(spl)); break;
    case 4: host_splitters<unsigned int>(host, n, xf, reinterpret_cast<unsigned int*>(spl)); break;
    case 8: host_splitters<unsigned long long>(host, n, xf, reinterpret_cast<unsigned long long*>(spl)); break;
    default: host_splitters<u128>(host, n, xf, reinterpret_cast<u128*>(spl)); break;
    }
    tr.mark("splitters chosen");
    SortCall pc;
    pc.ws = d_ws; pc.ws_bytes = ws_part; pc.st = cl.s;
    pc.flags = B200SORT_FLAG_NO_SKIP; pc.xf = xf; pc.splitters = spl; pc.nsplit = MAX_SPLITTERS;
    pc.hist_copy_words = PIPE_GROUPS;
    size_t bounce_issued = 0;
    int chunks = 0;
    for (int c = 0; c < CHUNKS; ++c) {
        const size_t off = (size_t)c * per_chunk;
        if (off >= n) break;
        const size_t cnt = std::min(per_chunk, n - off);
        if (staged) {  // the host thread feeds the bounce buffers; the split of chunk c runs under the copy of chunk c + 1
            CU(staged_copy_in(cl.bounce, d_in + off * kb, static_cast<const unsigned char*>(host) + off * kb, cnt * kb, cl.c,
                              &bounce_issued));
            CU(cudaEventCreateWithFlags(&landed[c], cudaEventDisableTiming));
            CU(cudaEventRecord(landed[c], cl.c));
        }
        CU(cudaStreamWaitEvent(cl.s, landed[c], 0));
        pc.elems = d_in + off * kb; pc.alt_out = d_out + off * kb; pc.n = cnt;
        pc.hist_copy = reinterpret_cast<unsigned long long*>(d_counts) + (size_t)c * PIPE_GROUPS;
        if ((rc = front_by_width<FAM_PARTITION>(kb, pc))) return rc;
        chunks = c + 1;
    }
    tr.mark("per-chunk range splits enqueued");
    if (tr.on) { cudaStreamSynchronize(cl.c); tr.mark("copy-in done"); }
    uint64_t counts[CHUNKS][PIPE_GROUPS];
    CU(cudaMemcpyAsync(counts, d_counts, (size_t)chunks * PIPE_GROUPS * 8, cudaMemcpyDeviceToHost, cl.s));
    CU(cudaStreamSynchronize(cl.s));
    tr.mark("range splits done, counts on the host");

    // range g is now 16 pieces, one per chunk slice of d_out (piece (c, g) starts where the
    // pieces (c, 0..g-1) end).  The input buffer is free: gather each range there, in range
    // order, and sort it in place -- the ranges end up contiguous and in order in d_in.
    size_t total[PIPE_GROUPS] = {}, max_g = 0;
    for (int g = 0; g < PIPE_GROUPS; ++g) {
        for (int c = 0; c < chunks; ++c) total[g] += (size_t)counts[c][g];
        max_g = std::max(max_g, total[g]);
    }
    const size_t need = b200sort_temp_bytes(max_g, kind);
    if (need > ws_sort_bytes) {  // skewed keys: one range holds far more than its share
        rc = pool_acquire(need, &cl.extra);
        if (rc) return rc;
        ws_sort = static_cast<unsigned char*>(cl.extra);
        ws_sort_bytes = need;
    }
    size_t piece_off[CHUNKS] = {};  // start of the next unread piece inside chunk c's slice
    size_t off = 0;
    for (int g = 0; g < PIPE_GROUPS; ++g) {  // all gathers + range sorts first: they are asynchronous
        if (total[g] == 0) continue;
        size_t at = off;
        for (int c = 0; c < chunks; ++c) {
            const size_t cnt = (size_t)counts[c][g];
            if (cnt == 0) continue;
            CU(cudaMemcpyAsync(d_in + at * kb, d_out + ((size_t)c * per_chunk + piece_off[c]) * kb, cnt * kb,
                               cudaMemcpyDeviceToDevice, cl.s));
            piece_off[c] += cnt;
            at += cnt;
        }
        rc = keys_device_dispatch(d_in + off * kb, ws_sort, ws_sort_bytes, total[g], kind, cl.s, 0, nullptr, 0);
        if (rc) return rc;
        CU(cudaEventCreateWithFlags(&cl.ev[g], cudaEventDisableTiming));
        CU(cudaEventRecord(cl.ev[g], cl.s));
        if (!staged) {  // pinned slice: the copy-out of a range is queued right behind its sort
            CU(cudaStreamWaitEvent(cl.c, cl.ev[g], 0));
            CU(cudaMemcpyAsync(static_cast<unsigned char*>(host) + off * kb, d_in + off * kb, total[g] * kb, cudaMemcpyDeviceToHost, cl.c));
        }
        off += total[g];
    }
    if (off != n) return fail(B200SORT_ERR_CUDA, "partition counts do not add up to n");
    tr.mark("range sorts enqueued");
    StagedCopyOut out(cl.bounce, cl.c);
    off = 0;
    for (int g = 0; g < PIPE_GROUPS && staged; ++g) {  // pageable slice: the host thread drains the bounce buffers, range by range
        const size_t cnt = total[g];
        if (cnt == 0) continue;
        CU(cudaStreamWaitEvent(cl.c, cl.ev[g], 0));
        CU(out.add(static_cast<unsigned char*>(host) + off * kb, d_in + off * kb, cnt * kb));
        off += cnt;
    }
    if (staged) CU(out.finish());
    CU(cudaStreamSynchronize(cl.s));
    tr.mark("range sorts done");
    CU(cudaStreamSynchronize(cl.c));
    tr.mark("copy-out done");
    return B200SORT_OK;
}

}  // namespace

// ---- workspace inside the library image (b200sort_keys_device_noalloc) ------------------
// Some hosts cannot hand us a d_tmp: they bind this library only (no cudaMalloc of their own), or they live
// in a process whose allocator owns all of HBM.  For them the K1/K2/K3 workspace is this array: part of the
// module, so it exists on every device from the moment the library's kernels are loaded there and costs
// B200SORT_NOALLOC_SCRATCH_BYTES of HBM per device until exit.
// In HOST code the name below is only a handle for the runtime, not a device address: handing it to a kernel
// would fault (or quietly hit host memory where the GPU can reach it).  The address a kernel may use comes
// from cudaGetSymbolAddress, asked once per device with that device current, and cached.
static __device__ __align__(256) unsigned char g_noalloc_scratch[B200SORT_NOALLOC_SCRATCH_BYTES];

namespace {
constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
void* g_noalloc_addr[NOALLOC_MAX_DEVICES] = {};  // device address of g_noalloc_scratch on device i; null = not looked up yet

// Device address of the scratch array on the calling thread's current device.
int noalloc_scratch_address(void** out) {
    int dev = -1;
    CU(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES)
        return fail(B200SORT_ERR_INVALID_ARG, "device ordinal " + std::to_string(dev) + " outside the static-scratch table");
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_addr[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);  // current device == dev
        if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(static scratch)");
        if (!p || ((uintptr_t)p % 256) != 0) return fail(B200SORT_ERR_CUDA, "static scratch has no usable device address");
        g_noalloc_addr[dev] = p;
    }
    *out = g_noalloc_addr[dev];
    return B200SORT_OK;
}
}  // namespace

// ---- stream-ordered workspace pool (b200sort_keys_device_pooled) --------------------------
// The reference's timer covers the allocation and release of its temp buffer (main.rs:32-34 around
// mod.rs:71-83, dropped at :176).  cudaMalloc + cudaFree around a 1 B-key sort cost 8.7 ms of 45.7
// (BENCH_r02 alloc_inside); a stream-ordered allocation from a pool that never gives memory back to the
// driver costs microseconds after the first call.  One pool per device, owned by the library (the
// process's default pool and its release threshold are left alone); b200sort_release_cached_memory trims it.
namespace {
std::mutex g_wspool_mu;
cudaMemPool_t g_wspool[NOALLOC_MAX_DEVICES] = {};

int workspace_pool(cudaMemPool_t* out) {
    int dev = -1;
    CU(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES)
        return fail(B200SORT_ERR_INVALID_ARG, "device ordinal " + std::to_string(dev) + " outside the workspace-pool table");
    std::lock_guard<std::mutex> lk(g_wspool_mu);
    if (!g_wspool[dev]) {
        cudaMemPoolProps props = {};
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = dev;
        cudaMemPool_t pool = nullptr;
        CU(cudaMemPoolCreate(&pool, &props));
        unsigned long long keep = ~0ull;  // freed blocks stay in the pool until it is trimmed
        const cudaError_t e = cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        if (e != cudaSuccess) {
            cudaMemPoolDestroy(pool);
            return cuda_fail(e, "cudaMemPoolSetAttribute(release threshold)");
        }
        g_wspool[dev] = pool;
    }
    *out = g_wspool[dev];
    return B200SORT_OK;
}

void workspace_pools_trim() {
    std::lock_guard<std::mutex> lk(g_wspool_mu);
    for (cudaMemPool_t pool : g_wspool)
        if (pool) cudaMemPoolTrimTo(pool, 0);
}
}  // namespace

// ---- exported C ABI ------------------------------------------------------------------
extern "C" {

int b200sort_version(void) { return B200SORT_VERSION; }

const char* b200sort_last_error(void) { return g_last_error.c_str(); }

size_t b200sort_key_bytes(int kind) {
    switch (kind) {
    case B200SORT_U8: case B200SORT_I8: return 1;
    case B200SORT_U16: case B200SORT_I16: return 2;
    case B200SORT_U32: case B200SORT_I32: case B200SORT_F32: return 4;
    case B200SORT_U64: case B200SORT_I64: case B200SORT_F64: return 8;
    case B200SORT_U128: case B200SORT_I128: return 16;
    default: return 0;
    }
}

size_t b200sort_temp_bytes(size_t n, int kind) {
    // the partition (K7) shares this workspace size: its tiles are those of the fallback variants
    return std::max(temp_by_width<FAM_KEYS>(b200sort_key_bytes(kind), n),
                    temp_by_width<FAM_PARTITION>(b200sort_key_bytes(kind), n));
}

int b200sort_keys_device(void* d_keys, void* d_tmp, size_t tmp_bytes, size_t n, int kind, void* stream) {
    return keys_device_dispatch(d_keys, d_tmp, tmp_bytes, n, kind, static_cast<cudaStream_t>(stream), 0, nullptr, 0);
}

int b200sort_keys_device_ex(void* d_keys, void* d_tmp, size_t tmp_bytes, size_t n, int kind, void* stream,
                            unsigned flags) {
    return keys_device_dispatch(d_keys, d_tmp, tmp_bytes, n, kind, static_cast<cudaStream_t>(stream), flags, nullptr, 0);
}

int b200sort_keys_device_timed(void* d_keys, void* d_tmp, size_t tmp_bytes, size_t n, int kind, void* stream,
                               unsigned flags, float* out_ms, int out_len) {
    if (!out_ms) return fail(B200SORT_ERR_INVALID_ARG, "out_ms is null");
    return keys_device_dispatch(d_keys, d_tmp, tmp_bytes, n, kind, static_cast<cudaStream_t>(stream), flags, out_ms, out_len);
}

// Device-resident sort with the workspace in g_noalloc_scratch: for callers that cannot allocate (see the
// header).  Same launch sequence as b200sort_keys_device -- only where d_tmp comes from differs -- and no
// cudaMalloc / cudaFree / allocating library call on this path.  The scratch is shared and dirty from the last
// call; the sort's own memset of its counters (front_def.cuh, do_zero) is what makes that harmless, and calls
// that overlap on the GPU race on it.
int b200sort_keys_device_noalloc(void* d_keys, size_t n, int kind, void* stream) {
    const size_t kb = b200sort_key_bytes(kind);
    if (kb == 0) return fail(B200SORT_ERR_UNSUPPORTED_KIND, "unknown key kind " + std::to_string(kind));
    if (n == 0) return B200SORT_OK;
    if (!d_keys) return fail(B200SORT_ERR_INVALID_ARG, "null device pointer");
    // refuse before touching the device: the layout grows with n, and n * kb alone may already overflow the array
    const size_t limit = B200SORT_NOALLOC_SCRATCH_BYTES;
    const size_t need = n > limit ? (size_t)-1 : b200sort_temp_bytes(n, kind);
    if (need > limit)
        return fail(B200SORT_ERR_INVALID_ARG,
                    "b200sort_keys_device_noalloc: " + std::to_string(n) + " keys of " + std::to_string(kb) + " bytes need " +
                        (n > limit ? std::string("more than ") + std::to_string(limit) : std::to_string(need)) +
                        " bytes of workspace, the built-in scratch holds " + std::to_string(limit) +
                        " (B200SORT_NOALLOC_SCRATCH_BYTES); use b200sort_keys_device with your own workspace");
    void* ws = nullptr;
    if (int rc = noalloc_scratch_address(&ws)) return rc;
    return keys_device_dispatch(d_keys, ws, limit, n, kind, static_cast<cudaStream_t>(stream), 0, nullptr, 0);
}

// Device-resident sort that takes and returns its workspace itself, in stream order (see the pool above):
// what a caller who would otherwise cudaMalloc / cudaFree around every b200sort_keys_device call should use.
int b200sort_keys_device_pooled(void* d_keys, size_t n, int kind, void* stream) {
    const size_t kb = b200sort_key_bytes(kind);
    if (kb == 0) return fail(B200SORT_ERR_UNSUPPORTED_KIND, "unknown key kind " + std::to_string(kind));
    if (n == 0) return B200SORT_OK;
    if (!d_keys) return fail(B200SORT_ERR_INVALID_ARG, "null device pointer");
    cudaMemPool_t pool = nullptr;
    if (int rc = workspace_pool(&pool)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t need = b200sort_temp_bytes(n, kind);
    void* ws = nullptr;
    const cudaError_t e = cudaMallocFromPoolAsync(&ws, need, pool, st);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return cuda_fail(e, "cudaMallocFromPoolAsync(workspace)");
    }
    const int rc = keys_device_dispatch(d_keys, ws, need, n, kind, st, 0, nullptr, 0);
    const cudaError_t e2 = cudaFreeAsync(ws, st);  // stream order: after the sort's last kernel
    if (rc) return rc;
    if (e2 != cudaSuccess) return cuda_fail(e2, "cudaFreeAsync(workspace)");
    return B200SORT_OK;
}

int b200sort_keys_host(void* keys, size_t n, int kind) {
    const size_t kb = b200sort_key_bytes(kind);
    if (kb == 0) return fail(B200SORT_ERR_UNSUPPORTED_KIND, "unknown key kind " + std::to_string(kind));
    if (n == 0) return B200SORT_OK;
    if (!keys) return fail(B200SORT_ERR_INVALID_ARG, "null host pointer");
    if (n * kb >= g_pipeline_min_bytes.load(std::memory_order_relaxed) && n >= 4096) return keys_host_pipelined(keys, n, kind);
    const size_t ws = b200sort_temp_bytes(n, kind);
    return host_roundtrip(keys, n * kb, ws, [&](void* d, void* w, size_t wb, cudaStream_t st) {
        return keys_device_dispatch(d, w, wb, n, kind, st, 0, nullptr, 0);
    });
}

int b200sort_release_cached_memory(void) {
    std::vector<PoolBlock> idle;
    {
        std::lock_guard<std::mutex> lk(g_pool_mu);
        for (int i = (int)g_pool.size() - 1; i >= 0; --i)
            if (!g_pool[i].busy) {
                idle.push_back(g_pool[i]);
                g_pool.erase(g_pool.begin() + i);
            }
    }
    int prev = -1;
    cudaGetDevice(&prev);
    for (auto& b : idle) {
        cudaSetDevice(b.dev);
        cudaFree(b.p);
    }
    if (prev >= 0) cudaSetDevice(prev);
    workspace_pools_trim();  // blocks b200sort_keys_device_pooled has returned (work still in flight keeps its own)
    cudaGetLastError();
    return B200SORT_OK;
}

int b200sort_set_host_pipeline_min_bytes(size_t bytes) {
    g_pipeline_min_bytes.store(bytes, std::memory_order_relaxed);
    return B200SORT_OK;
}

size_t b200sort_records_temp_bytes(size_t n, size_t rec_bytes, int kind) {
    (void)kind;
    if (rec_bytes != 8 && rec_bytes != 16) return 0;
    return temp_by_width<FAM_RECORDS>(rec_bytes, n);
}

int b200sort_records_device(void* d_recs, void* d_tmp, size_t tmp_bytes, size_t n, size_t rec_bytes,
                            size_t key_off, int kind, void* stream) {
    const int rc = check_record_shape(rec_bytes, key_off, kind);
    if (rc) return rc;
    SortCall c;
    c.elems = d_recs; c.ws = d_tmp; c.ws_bytes = tmp_bytes; c.n = n; c.st = static_cast<cudaStream_t>(stream);
    c.xf = kind_xf(kind); c.key_shift = (int)key_off * 8;
    return front_by_width<FAM_RECORDS>(rec_bytes, c);
}

int b200sort_records_host(void* recs, size_t n, size_t rec_bytes, size_t key_off, int kind) {
    const int rc = check_record_shape(rec_bytes, key_off, kind);
    if (rc) return rc;
    if (n == 0) return B200SORT_OK;
    if (!recs) return fail(B200SORT_ERR_INVALID_ARG, "null host pointer");
    const size_t ws = b200sort_records_temp_bytes(n, rec_bytes, kind);
    return host_roundtrip(recs, n * rec_bytes, ws, [&](void* d, void* w, size_t wb, cudaStream_t st) {
        return b200sort_records_device(d, w, wb, n, rec_bytes, key_off, kind, st);
    });
}

int b200sort_partition_device(const void* d_in, void* d_out, void* d_tmp, size_t tmp_bytes, size_t n, int kind,
                              const void* splitters, int nsplit, uint64_t* d_counts, void* stream) {
    if (nsplit < 0 || nsplit > MAX_SPLITTERS) return fail(B200SORT_ERR_INVALID_ARG, "nsplit must be in [0, 15]");
    if (nsplit > 0 && !splitters) return fail(B200SORT_ERR_INVALID_ARG, "null splitters");
    if (!d_counts || (n && (!d_in || !d_out))) return fail(B200SORT_ERR_INVALID_ARG, "null device pointer");
    const size_t kb = b200sort_key_bytes(kind);
    if (kb == 0) return fail(B200SORT_ERR_UNSUPPORTED_KIND, "unknown key kind " + std::to_string(kind));
    SortCall c;
    c.elems = const_cast<void*>(d_in); c.alt_out = d_out; c.ws = d_tmp; c.ws_bytes = tmp_bytes; c.n = n;
    c.st = static_cast<cudaStream_t>(stream); c.flags = B200SORT_FLAG_NO_SKIP; c.xf = kind_xf(kind);
    c.hist_copy = reinterpret_cast<unsigned long long*>(d_counts); c.hist_copy_words = nsplit + 1;
    c.splitters = splitters; c.nsplit = nsplit;
    return front_by_width<FAM_PARTITION>(kb, c);
}

static int partition_phase(int phase, const void* d_in, void* d_tmp, size_t tmp_bytes, size_t n, int kind,
                           const void* splitters, int nsplit, uint64_t* d_counts, void* const* peer_bases,
                           const uint64_t* d_peer_offsets, void* stream, unsigned bucket_mask = 0xFFFFFFFFu,
                           bool rezero = false) {
    if (nsplit < 0 || nsplit > MAX_SPLITTERS) return fail(B200SORT_ERR_INVALID_ARG, "nsplit must be in [0, 15]");
    if (nsplit > 0 && !splitters) return fail(B200SORT_ERR_INVALID_ARG, "null splitters");
    if (n && !d_in) return fail(B200SORT_ERR_INVALID_ARG, "null device pointer");
    if (phase == SortCall::PHASE_COUNT && !d_counts) return fail(B200SORT_ERR_INVALID_ARG, "null d_counts");
    if (phase == SortCall::PHASE_SCATTER && (!peer_bases || !d_peer_offsets))
        return fail(B200SORT_ERR_INVALID_ARG, "null peer_bases / d_peer_offsets");
    const size_t kb = b200sort_key_bytes(kind);
    if (kb == 0) return fail(B200SORT_ERR_UNSUPPORTED_KIND, "unknown key kind " + std::to_string(kind));
    SortCall c;
    c.elems = const_cast<void*>(d_in); c.ws = d_tmp; c.ws_bytes = tmp_bytes; c.n = n;
    // every bucket has its own base (peer_bases); alt_out only has to be non-null so
    // that the front treats the call as single-pass with no copy-back
    c.alt_out = phase == SortCall::PHASE_SCATTER ? peer_bases[0] : const_cast<void*>(d_in);
    c.st = static_cast<cudaStream_t>(stream); c.flags = B200SORT_FLAG_NO_SKIP; c.xf = kind_xf(kind);
    if (phase == SortCall::PHASE_COUNT) {
        c.hist_copy = reinterpret_cast<unsigned long long*>(d_counts); c.hist_copy_words = nsplit + 1;
    }
    c.splitters = splitters; c.nsplit = nsplit; c.phase = phase;
    c.bucket_mask = bucket_mask; c.rezero = rezero;
    c.peer_bases = peer_bases; c.peer_offsets = reinterpret_cast<const unsigned long long*>(d_peer_offsets);
    return front_by_width<FAM_PARTITION>(kb, c);
}

int b200sort_partition_count_device(const void* d_in, void* d_tmp, size_t tmp_bytes, size_t n, int kind,
                                    const void* splitters, int nsplit, uint64_t* d_counts, void* stream) {
    return partition_phase(SortCall::PHASE_COUNT, d_in, d_tmp, tmp_bytes, n, kind, splitters, nsplit, d_counts, nullptr,
                           nullptr, stream);
}

int b200sort_partition_exchange_device(const void* d_in, void* d_tmp, size_t tmp_bytes, size_t n, int kind,
                                       const void* splitters, int nsplit, void* const* peer_bases,
                                       const uint64_t* d_peer_offsets, void* stream) {
    return partition_phase(SortCall::PHASE_SCATTER, d_in, d_tmp, tmp_bytes, n, kind, splitters, nsplit, nullptr,
                           peer_bases, d_peer_offsets, stream);
}

int b200sort_partition_claim_device(const void* d_in, void* d_tmp, size_t tmp_bytes, size_t n, int kind, const void* splitters,
                                    int nsplit, void* const* region_bases, const uint64_t* cursor_addrs,
                                    const uint64_t* region_caps, uint32_t* d_flags, void* stream) {
    if (nsplit < 0 || nsplit > MAX_SPLITTERS) return fail(B200SORT_ERR_INVALID_ARG, "nsplit must be in [0, 15]");
    if (nsplit > 0 && !splitters) return fail(B200SORT_ERR_INVALID_ARG, "null splitters");
    if (n && !d_in) return fail(B200SORT_ERR_INVALID_ARG, "null device pointer");
    if (!region_bases || !cursor_addrs || !region_caps || !d_flags) return fail(B200SORT_ERR_INVALID_ARG, "null claim argument");
    const size_t kb = b200sort_key_bytes(kind);
    if (kb == 0) return fail(B200SORT_ERR_UNSUPPORTED_KIND, "unknown key kind " + std::to_string(kind));
    SortCall c;
    c.elems = const_cast<void*>(d_in); c.ws = d_tmp; c.ws_bytes = tmp_bytes; c.n = n;
    c.alt_out = region_bases[0];
    c.st = static_cast<cudaStream_t>(stream); c.flags = B200SORT_FLAG_NO_SKIP; c.xf = kind_xf(kind);
    c.splitters = splitters; c.nsplit = nsplit; c.phase = SortCall::PHASE_CLAIM;
    c.peer_bases = region_bases;
    c.claim_cursors = reinterpret_cast<const unsigned long long*>(cursor_addrs);
    c.claim_caps = reinterpret_cast<const unsigned long long*>(region_caps);
    c.claim_flags = d_flags;
    return front_by_width<FAM_PARTITION>(kb, c);
}

int b200sort_partition_exchange_part_device(const void* d_in, void* d_tmp, size_t tmp_bytes, size_t n, int kind,
                                            const void* splitters, int nsplit, void* const* peer_bases,
                                            const uint64_t* d_peer_offsets, unsigned bucket_mask, int again, void* stream) {
    return partition_phase(SortCall::PHASE_SCATTER, d_in, d_tmp, tmp_bytes, n, kind, splitters, nsplit, nullptr,
                           peer_bases, d_peer_offsets, stream, bucket_mask, again != 0);
}

unsigned long long b200sort_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

void b200sort_debug_small_sort(int on) { g_small_sort.store(on ? 1 : 0, std::memory_order_relaxed); }

void b200sort_debug_narrow_fill(int on) { g_narrow_fill.store(on ? 1 : 0, std::memory_order_relaxed); }

void b200sort_debug_force_rerank(int on) { g_force_rerank.store(on ? 1 : 0, std::memory_order_relaxed); }

int b200sort_debug_reranked_tiles(void* d_tmp, size_t tmp_bytes, size_t n, int kind, size_t rec_bytes, void* stream,
                                  unsigned* out) {
    if (!d_tmp || !out) return fail(B200SORT_ERR_INVALID_ARG, "null pointer");
    *out = 0;
    if (n == 0) return B200SORT_OK;
    const size_t kb = b200sort_key_bytes(kind);
    if (kb == 0) return fail(B200SORT_ERR_UNSUPPORTED_KIND, "unknown key kind " + std::to_string(kind));
    SortCall c;
    c.elems = d_tmp;  // not touched; only has to be non-null
    c.ws = d_tmp; c.ws_bytes = tmp_bytes; c.n = n; c.st = static_cast<cudaStream_t>(stream);
    c.phase = SortCall::PHASE_READ_STATS; c.out_stats = out;
    return rec_bytes ? front_by_width<FAM_RECORDS>(rec_bytes, c) : front_by_width<FAM_KEYS>(kb, c);
}

int b200sort_set_variant(int key_bytes, int variant) {
    // key_bytes selects the key-only table of that width; 108 / 116 select the
    // record tables (8- and 16-byte records)
    int fam = FAM_KEYS, w = key_bytes, nv = 0;
    if (key_bytes == 108 || key_bytes == 116) { fam = FAM_RECORDS; w = key_bytes - 100; }
    if (fam == FAM_KEYS) {
        switch (w) {
        case 1: nv = num_variants<FAM_KEYS, 1>(); break;
        case 2: nv = num_variants<FAM_KEYS, 2>(); break;
        case 4: nv = num_variants<FAM_KEYS, 4>(); break;
        case 8: nv = num_variants<FAM_KEYS, 8>(); break;
        case 16: nv = num_variants<FAM_KEYS, 16>(); break;
        default: return fail(B200SORT_ERR_INVALID_ARG, "key_bytes must be 1, 2, 4, 8, 16 (keys) or 108, 116 (records)");
        }
    } else {
        nv = w == 8 ? num_variants<FAM_RECORDS, 8>() : num_variants<FAM_RECORDS, 16>();
    }
    if (variant < -1 || variant >= nv) return fail(B200SORT_ERR_INVALID_ARG, "no such variant");
    g_variant[fam][w].store(variant + 1, std::memory_order_relaxed);
    return B200SORT_OK;
}

int b200sort_num_variants(int key_bytes) {
    switch (key_bytes) {
    case 1: return num_variants<FAM_KEYS, 1>();
    case 2: return num_variants<FAM_KEYS, 2>();
    case 4: return num_variants<FAM_KEYS, 4>();
    case 8: return num_variants<FAM_KEYS, 8>();
    case 16: return num_variants<FAM_KEYS, 16>();
    case 108: return num_variants<FAM_RECORDS, 8>();
    case 116: return num_variants<FAM_RECORDS, 16>();
    default: return 0;
    }
}

}  // extern "C"

// ---- all GPUs of one process behind one call (b200sort_mgpu_*) ----------
// The caller of `slice.radix_sort()` is ONE process; the rank-per-process path
// (radix_sort_b200/mgpu.py) needs a launcher around it.  This is the same algorithm
// -- sampled splitters, K7 count, K7 scatter fused with the exchange (every key is
// stored straight into its destination GPU's receive buffer over NVLink), local
// sorts -- driven by one host thread over the devices of the communicator.  What
// the ranks exchange through all_gather there is a host array here, and a barrier
// is a loop of cudaStreamSynchronize.
struct b200sort_comm {
    int ndev = 0;
    int dev[MAX_SPLITTERS + 1] = {};
    cudaStream_t st[MAX_SPLITTERS + 1] = {};
    void* recv[MAX_SPLITTERS + 1] = {};   // receive buffer = result of the device
    void* part[MAX_SPLITTERS + 1] = {};   // K7 state + counts + offsets + samples
    void* sortws[MAX_SPLITTERS + 1] = {};  // workspace of the local sort
    void* in[MAX_SPLITTERS + 1] = {};      // b200sort_mgpu_keys_host: the device's part of the host slice
    size_t recv_bytes[MAX_SPLITTERS + 1] = {}, part_bytes[MAX_SPLITTERS + 1] = {}, sortws_bytes[MAX_SPLITTERS + 1] = {},
           in_bytes[MAX_SPLITTERS + 1] = {};
    uint64_t* h_counts = nullptr;  // pinned: a device->host copy into pageable memory blocks the host until it is done,
                                   // which would run the devices' count phases one after the other
    std::mutex mu;  // one sort at a time per communicator
};

namespace {

constexpr size_t MGPU_SAMPLES = (size_t)1 << 14;  // per device, scaled down for smaller shards
std::atomic<size_t> g_mgpu_host_min_bytes{(size_t)32 << 20};  // smaller host slices stay on the first device

struct DeviceGuard {
    int prev = -1;
    DeviceGuard() { cudaGetDevice(&prev); }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

template <typename U>
__global__ void gather_samples_kernel(const U* __restrict__ in, size_t stride, unsigned long long salt, U* __restrict__ out,
                                      unsigned samples) {
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < samples) out[i] = in[(size_t)i * stride + (size_t)(mix64(salt ^ mix64(i)) % stride)];
}

int launch_gather(size_t kb, const void* in, size_t stride, unsigned long long salt, void* out, unsigned samples, cudaStream_t st) {
    const unsigned blocks = (samples + 255) / 256;
    switch (kb) {
    case 1: gather_samples_kernel<<<blocks, 256, 0, st>>>(static_cast<const unsigned char*>(in), stride, salt, static_cast<unsigned char*>(out), samples); break;
    case 2: gather_samples_kernel<<<blocks, 256, 0, st>>>(static_cast<const unsigned short*>(in), stride, salt, static_cast<unsigned short*>(out), samples); break;
    case 4: gather_samples_kernel<<<blocks, 256, 0, st>>>(static_cast<const unsigned int*>(in), stride, salt, static_cast<unsigned int*>(out), samples); break;
    case 8: gather_samples_kernel<<<blocks, 256, 0, st>>>(static_cast<const unsigned long long*>(in), stride, salt, static_cast<unsigned long long*>(out), samples); break;
    default: gather_samples_kernel<<<blocks, 256, 0, st>>>(static_cast<const u128*>(in), stride, salt, static_cast<u128*>(out), samples); break;
    }
    g_launches.fetch_add(1, std::memory_order_relaxed);
    CU(cudaGetLastError());
    return B200SORT_OK;
}

// device memory of the communicator grows on demand and is kept between sorts
int grow(void** p, size_t* have, size_t need) {
    if (*have >= need && *p) return B200SORT_OK;
    if (*p) cudaFree(*p);
    *p = nullptr; *have = 0;
    need = align_up(need + need / 16 + 256, 256);
    const cudaError_t e = cudaMalloc(p, need);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(B200SORT_ERR_OOM, std::string("cudaMalloc of ") + std::to_string(need) + " bytes: " + cudaGetErrorString(e));
    }
    *have = need;
    return B200SORT_OK;
}

int sync_all(b200sort_comm* c) {
    for (int r = 0; r < c->ndev; ++r) {
        CU(cudaSetDevice(c->dev[r]));
        CU(cudaStreamSynchronize(c->st[r]));
    }
    return B200SORT_OK;
}

}  // namespace

extern "C" {

int b200sort_mgpu_create(b200sort_comm** out, int ndev, const int* devs) {
    if (!out) return fail(B200SORT_ERR_INVALID_ARG, "null communicator pointer");
    *out = nullptr;
    if (ndev < 1 || ndev > MAX_SPLITTERS + 1) return fail(B200SORT_ERR_INVALID_ARG, "ndev must be in [1, 16]");
    int have = 0;
    CU(cudaGetDeviceCount(&have));
    DeviceGuard guard;
    b200sort_comm* c = new b200sort_comm;
    c->ndev = ndev;
    for (int r = 0; r < ndev; ++r) {
        c->dev[r] = devs ? devs[r] : r;
        if (c->dev[r] < 0 || c->dev[r] >= have) {
            delete c;
            return fail(B200SORT_ERR_INVALID_ARG, "no such CUDA device");
        }
    }
    int rc = B200SORT_OK;
    for (int r = 0; r < ndev && rc == B200SORT_OK; ++r) {
        cudaError_t e = cudaSetDevice(c->dev[r]);
        for (int q = 0; q < ndev && e == cudaSuccess; ++q) {  // a device may be named twice (two shards on one GPU)
            if (c->dev[q] == c->dev[r]) continue;
            int can = 0;
            e = cudaDeviceCanAccessPeer(&can, c->dev[r], c->dev[q]);
            if (e == cudaSuccess && !can) {
                rc = fail(B200SORT_ERR_CUDA, "no peer access between devices " + std::to_string(c->dev[r]) + " and " + std::to_string(c->dev[q]));
                break;
            }
            if (e == cudaSuccess) {
                e = cudaDeviceEnablePeerAccess(c->dev[q], 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
            }
        }
        if (rc == B200SORT_OK && e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->st[r], cudaStreamNonBlocking);
        if (rc == B200SORT_OK && e != cudaSuccess) rc = cuda_fail(e, "b200sort_mgpu_create");
    }
    if (rc == B200SORT_OK) {
        const cudaError_t e = cudaHostAlloc(reinterpret_cast<void**>(&c->h_counts),
                                            sizeof(uint64_t) * (MAX_SPLITTERS + 1) * (MAX_SPLITTERS + 1), cudaHostAllocPortable);
        if (e != cudaSuccess) rc = cuda_fail(e, "cudaHostAlloc(count matrix)");
    }
    if (rc != B200SORT_OK) {
        b200sort_mgpu_destroy(c);
        return rc;
    }
    *out = c;
    return B200SORT_OK;
}

int b200sort_mgpu_destroy(b200sort_comm* c) {
    if (!c) return B200SORT_OK;
    DeviceGuard guard;
    for (int r = 0; r < c->ndev; ++r) {
        if (cudaSetDevice(c->dev[r]) != cudaSuccess) continue;
        if (c->st[r]) { cudaStreamSynchronize(c->st[r]); cudaStreamDestroy(c->st[r]); }
        if (c->recv[r]) cudaFree(c->recv[r]);
        if (c->part[r]) cudaFree(c->part[r]);
        if (c->sortws[r]) cudaFree(c->sortws[r]);
        if (c->in[r]) cudaFree(c->in[r]);
    }
    if (c->h_counts) cudaFreeHost(c->h_counts);
    cudaGetLastError();
    delete c;
    return B200SORT_OK;
}

// the sort proper; the caller holds c->mu and restores the current device
static int mgpu_sort_locked(b200sort_comm* c, const void* const* d_shards, const size_t* n_in, int kind, void** d_out,
                            size_t* n_out) {
    const size_t kb = b200sort_key_bytes(kind);
    const int G = c->ndev;
    Trace tr;
    int rc;
    size_t max_n = 0;
    for (int r = 0; r < G; ++r) {
        if (n_in[r] && !d_shards[r]) return fail(B200SORT_ERR_INVALID_ARG, "null shard");
        max_n = std::max(max_n, n_in[r]);
    }
    // layout of part[r]: K7 state | counts (G u64) | offsets (G u64) | pooled samples | workspace of their sort
    // (the last two are used on the first device only)
    const size_t pool_cap = align_up((size_t)G * MGPU_SAMPLES * 16, 256);
    const size_t pool_ws = align_up(b200sort_temp_bytes((size_t)G * MGPU_SAMPLES, kind), 256);
    size_t state_bytes[MAX_SPLITTERS + 1], samples[MAX_SPLITTERS + 1], sample_base[MAX_SPLITTERS + 2] = {0};
    for (int r = 0; r < G; ++r) {
        state_bytes[r] = align_up(temp_by_width<FAM_PARTITION>(kb, n_in[r]), 256);
        // sample count proportional to the shard, so that the pooled quantiles weigh every key alike
        samples[r] = n_in[r] ? std::max<size_t>(1, std::min(n_in[r], (size_t)((double)MGPU_SAMPLES * (double)n_in[r] / (double)max_n))) : 0;
        sample_base[r + 1] = sample_base[r] + samples[r];
        CU(cudaSetDevice(c->dev[r]));
        if ((rc = grow(&c->part[r], &c->part_bytes[r], state_bytes[r] + 512 + (r == 0 ? pool_cap + pool_ws : 0)))) return rc;
    }
    auto counts_of = [&](int r) { return reinterpret_cast<uint64_t*>(static_cast<unsigned char*>(c->part[r]) + state_bytes[r]); };
    auto offsets_of = [&](int r) { return counts_of(r) + 32; };
    unsigned char* pool = static_cast<unsigned char*>(c->part[0]) + state_bytes[0] + 512;
    const size_t pooled = sample_base[G];
    for (int r = 0; r < G; ++r) { d_out[r] = nullptr; n_out[r] = 0; }
    if (pooled == 0) return B200SORT_OK;  // nothing to sort

    // 1. splitters: every device stores samples of its shard straight into the first device's
    //    pool (peer stores); the pool is sorted there with the library's own sort (host
    //    std::sort of 32 K keys took 1.5 ms) and its quantiles come back to the host
    for (int r = 0; r < G; ++r) {
        if (!samples[r]) continue;
        CU(cudaSetDevice(c->dev[r]));
        if ((rc = launch_gather(kb, d_shards[r], n_in[r] / samples[r], 0xB200ull + (unsigned long long)r, pool + sample_base[r] * kb,
                                (unsigned)samples[r], c->st[r])))
            return rc;
    }
    if ((rc = sync_all(c))) return rc;
    CU(cudaSetDevice(c->dev[0]));
    if ((rc = keys_device_dispatch(pool, pool + pool_cap, pool_ws, pooled, kind, c->st[0], 0, nullptr, 0))) return rc;
    alignas(16) unsigned char spl[MAX_SPLITTERS * 16];
    for (int j = 0; j + 1 < G; ++j)
        CU(cudaMemcpyAsync(spl + (size_t)j * kb, pool + std::min(pooled - 1, pooled * (size_t)(j + 1) / G) * kb, kb, cudaMemcpyDeviceToHost, c->st[0]));
    CU(cudaStreamSynchronize(c->st[0]));
    tr.mark("mgpu: splitters chosen");

    // 2. K7 count on every device; the G x G count matrix comes to the host
    uint64_t cnt[MAX_SPLITTERS + 1][MAX_SPLITTERS + 1] = {};
    for (int r = 0; r < G; ++r) {
        CU(cudaSetDevice(c->dev[r]));
        if ((rc = partition_phase(SortCall::PHASE_COUNT, d_shards[r], c->part[r], state_bytes[r], n_in[r], kind, spl, G - 1,
                                  counts_of(r), nullptr, nullptr, c->st[r])))
            return rc;
        CU(cudaMemcpyAsync(c->h_counts + (size_t)r * (MAX_SPLITTERS + 1), counts_of(r), (size_t)G * 8, cudaMemcpyDeviceToHost, c->st[r]));
    }
    if ((rc = sync_all(c))) return rc;
    for (int r = 0; r < G; ++r)
        for (int j = 0; j < G; ++j) cnt[r][j] = c->h_counts[(size_t)r * (MAX_SPLITTERS + 1) + j];
    tr.mark("mgpu: counts on the host");

    // 3. receive buffers and the offset at which every sender's keys start in them
    uint64_t off[MAX_SPLITTERS + 1][MAX_SPLITTERS + 1] = {};
    size_t total_in = 0, total_out = 0;
    for (int j = 0; j < G; ++j) {
        size_t at = 0;
        for (int r = 0; r < G; ++r) { off[r][j] = at; at += (size_t)cnt[r][j]; }
        n_out[j] = at;
        total_out += at;
        total_in += n_in[j];
        CU(cudaSetDevice(c->dev[j]));
        if ((rc = grow(&c->recv[j], &c->recv_bytes[j], at * kb))) return rc;
        if ((rc = grow(&c->sortws[j], &c->sortws_bytes[j], b200sort_temp_bytes(at, kind)))) return rc;
    }
    if (total_in != total_out) return fail(B200SORT_ERR_CUDA, "partition counts do not add up to the input");

    // 4. K7 scatter: every key goes straight to its destination's receive buffer
    for (int r = 0; r < G; ++r) {
        if (!n_in[r]) continue;
        CU(cudaSetDevice(c->dev[r]));
        CU(cudaMemcpyAsync(offsets_of(r), off[r], (size_t)G * 8, cudaMemcpyHostToDevice, c->st[r]));
        if ((rc = partition_phase(SortCall::PHASE_SCATTER, d_shards[r], c->part[r], state_bytes[r], n_in[r], kind, spl, G - 1,
                                  nullptr, c->recv, offsets_of(r), c->st[r])))
            return rc;
    }
    if ((rc = sync_all(c))) return rc;  // every receive buffer is complete
    tr.mark("mgpu: keys exchanged");

    // 5. local sorts, all devices at once
    for (int j = 0; j < G; ++j) {
        d_out[j] = c->recv[j];
        if (!n_out[j]) continue;
        CU(cudaSetDevice(c->dev[j]));
        if ((rc = keys_device_dispatch(c->recv[j], c->sortws[j], c->sortws_bytes[j], n_out[j], kind, c->st[j], 0, nullptr, 0))) return rc;
    }
    rc = sync_all(c);
    tr.mark("mgpu: local sorts done");
    return rc;
}

int b200sort_mgpu_keys_device(b200sort_comm* c, const void* const* d_shards, const size_t* n_in, int kind,
                              void** d_out, size_t* n_out) {
    if (!c || !d_shards || !n_in || !d_out || !n_out) return fail(B200SORT_ERR_INVALID_ARG, "null argument");
    if (b200sort_key_bytes(kind) == 0) return fail(B200SORT_ERR_UNSUPPORTED_KIND, "unknown key kind " + std::to_string(kind));
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard guard;
    return mgpu_sort_locked(c, d_shards, n_in, kind, d_out, n_out);
}

// one host slice over all devices of the communicator: every device copies its part
// in over its own PCIe link, the parts are sorted as one key set, every device copies
// its range of the result back to where it belongs in the slice
int b200sort_mgpu_keys_host(b200sort_comm* c, void* keys, size_t n, int kind) {
    if (!c) return fail(B200SORT_ERR_INVALID_ARG, "null communicator");
    const size_t kb = b200sort_key_bytes(kind);
    if (kb == 0) return fail(B200SORT_ERR_UNSUPPORTED_KIND, "unknown key kind " + std::to_string(kind));
    if (n == 0) return B200SORT_OK;
    if (!keys) return fail(B200SORT_ERR_INVALID_ARG, "null keys");
    const int G = c->ndev;
    DeviceGuard guard;
    // a pageable slice (a plain Vec<T>) moves at the speed of the host-side staging copy
    // (host_staging.cuh), and the driver's own staging, which G plain copies would use, is
    // far slower still (2 GPUs, 1 B u64: 733 ms against 310 on one device): it stays on the
    // first device until the bounce buffers are per device
    if (G == 1 || n * kb < g_mgpu_host_min_bytes.load(std::memory_order_relaxed) || is_pageable(keys)) {
        CU(cudaSetDevice(c->dev[0]));
        return b200sort_keys_host(keys, n, kind);
    }
    std::lock_guard<std::mutex> lk(c->mu);
    Trace tr;
    int rc;
    unsigned char* host = static_cast<unsigned char*>(keys);
    size_t n_in[MAX_SPLITTERS + 1] = {}, n_out[MAX_SPLITTERS + 1] = {}, at[MAX_SPLITTERS + 2] = {0};
    void* d_out[MAX_SPLITTERS + 1] = {};
    const size_t per = (n + G - 1) / G;
    for (int r = 0; r < G; ++r) {
        const size_t lo = std::min(n, (size_t)r * per);
        n_in[r] = std::min(per, n - lo);
        at[r + 1] = lo + n_in[r];
        CU(cudaSetDevice(c->dev[r]));
        if ((rc = grow(&c->in[r], &c->in_bytes[r], n_in[r] * kb))) return rc;
    }
    // one host thread per device for the copies (the DMAs of a pinned slice run side by side)
    auto copy_all = [&](bool in_dir, const size_t* offs, const size_t* counts) -> int {
        cudaError_t err[MAX_SPLITTERS + 1] = {};
        std::vector<std::thread> th;
        for (int r = 0; r < G; ++r)
            th.emplace_back([&, r] {
                if (!counts[r]) return;
                cudaError_t e = cudaSetDevice(c->dev[r]);
                if (e == cudaSuccess)
                    e = in_dir ? cudaMemcpyAsync(c->in[r], host + offs[r] * kb, counts[r] * kb, cudaMemcpyHostToDevice, c->st[r])
                               : cudaMemcpyAsync(host + offs[r] * kb, c->recv[r], counts[r] * kb, cudaMemcpyDeviceToHost, c->st[r]);
                if (e == cudaSuccess) e = cudaStreamSynchronize(c->st[r]);
                err[r] = e;
            });
        for (auto& t : th) t.join();
        for (int r = 0; r < G; ++r)
            if (err[r] != cudaSuccess) return cuda_fail(err[r], in_dir ? "copy-in of a device's part" : "copy-out of a device's range");
        return B200SORT_OK;
    };
    if ((rc = copy_all(true, at, n_in))) return rc;
    tr.mark("mgpu host: parts copied in");
    if ((rc = mgpu_sort_locked(c, c->in, n_in, kind, d_out, n_out))) return rc;
    size_t out_at[MAX_SPLITTERS + 2] = {0};
    for (int r = 0; r < G; ++r) out_at[r + 1] = out_at[r] + n_out[r];
    if (out_at[G] != n) return fail(B200SORT_ERR_CUDA, "ranges do not add up to n");
    rc = copy_all(false, out_at, n_out);
    tr.mark("mgpu host: ranges copied out");
    return rc;
}

int b200sort_mgpu_set_host_min_bytes(size_t bytes) {
    g_mgpu_host_min_bytes.store(bytes, std::memory_order_relaxed);
    return B200SORT_OK;
}

}  // extern "C"
