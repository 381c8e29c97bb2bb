// b200sort_api.cu -- the C ABI of libb200sort.so (include/b200sort.h): dispatch
// over key kinds, workspace layout, launch sequence of K1..K3, host-slice path.
//
// Replaces RadixSort::radix_sort for [T] (reference src/radix_sort/mod.rs:61-176).
// There is deliberately no CPU path in this file: every compute entry point
// either launches the sm_100a kernels or returns an error.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/b200sort.h"
#include "sort_kernels.cuh"

using namespace b200;

namespace {

thread_local std::string g_last_error;
std::atomic<unsigned long long> g_launches{0};
std::atomic<int> g_variant[17];  // indexed by element bytes; 0 = default

int fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}
int cuda_fail(cudaError_t e, const char* what) {
    return fail(e == cudaErrorMemoryAllocation ? B200SORT_ERR_OOM : B200SORT_ERR_CUDA,
                std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")");
}
#define CU(call)                                              \
    do {                                                      \
        cudaError_t e_ = (call);                              \
        if (e_ != cudaSuccess) return cuda_fail(e_, #call);   \
    } while (0)

}  // namespace
namespace b200 { void set_last_error(const std::string& msg) { g_last_error = msg; } }  // used by aux_kernels.cu
namespace {

size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---- kernel variants (THREADS x ITEMS per element width) ----------------------
struct VariantDesc { int threads, items, min_blocks; };

// The table a variant index selects from, per element width.  Index 0 is the
// default; the others exist for on-GPU sweeps (b200sort_set_variant).
template <int EB> struct Variants;
template <> struct Variants<1>  { static constexpr VariantDesc v[] = {{512, 32, 2}, {512, 16, 3}, {256, 32, 4}}; };
template <> struct Variants<2>  { static constexpr VariantDesc v[] = {{512, 24, 2}, {512, 16, 3}, {256, 24, 4}}; };
template <> struct Variants<4>  { static constexpr VariantDesc v[] = {{512, 16, 2}, {512, 22, 2}, {512, 16, 3}, {512, 12, 4}, {256, 16, 6}, {256, 24, 4}, {384, 20, 3}, {1024, 12, 1}}; };
template <> struct Variants<8>  { static constexpr VariantDesc v[] = {{512, 12, 2}, {512, 16, 2}, {512, 10, 3}, {512, 8, 4}, {256, 12, 6}, {256, 16, 4}, {384, 12, 3}, {1024, 8, 1}}; };
template <> struct Variants<16> { static constexpr VariantDesc v[] = {{256, 12, 2}, {512, 8, 2}, {256, 8, 4}, {512, 6, 2}}; };

template <int EB> constexpr int num_variants() { return (int)(sizeof(Variants<EB>::v) / sizeof(VariantDesc)); }
template <int EB> int min_tile() {
    int m = 1 << 30;
    for (int i = 0; i < num_variants<EB>(); ++i) {
        const int t = Variants<EB>::v[i].threads * Variants<EB>::v[i].items;
        if (t < m) m = t;
    }
    return m;
}

template <typename Tr, int THREADS, int ITEMS, int MINB>
cudaError_t launch_pass_t(const PassArgs<Tr>& a, cudaStream_t st) {
    using SM = PassSmem<Tr, THREADS, ITEMS>;
    auto kern = onesweep_pass_kernel<Tr, THREADS, ITEMS, MINB>;
    static std::atomic<unsigned> configured{0};  // bit per device
    int dev = 0;
    cudaGetDevice(&dev);
    if (!(configured.load(std::memory_order_acquire) & (1u << (dev & 31)))) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SM::TOTAL);
        if (e != cudaSuccess) return e;
        configured.fetch_or(1u << (dev & 31), std::memory_order_release);
    }
    kern<<<a.num_tiles, THREADS, SM::TOTAL, st>>>(a);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return cudaGetLastError();
}

template <typename Tr, int EB, int I>
cudaError_t launch_pass_idx(int variant, const PassArgs<Tr>& a, cudaStream_t st) {
    if constexpr (I >= num_variants<EB>()) {
        return cudaErrorInvalidValue;
    } else {
        if (variant == I) {
            constexpr VariantDesc d = Variants<EB>::v[I];
            return launch_pass_t<Tr, d.threads, d.items, d.min_blocks>(a, st);
        }
        return launch_pass_idx<Tr, EB, I + 1>(variant, a, st);
    }
}

template <int EB> int current_variant() {
    int v = g_variant[EB].load(std::memory_order_relaxed);
    if (v < 0 || v >= num_variants<EB>()) v = 0;
    return v;
}

// ---- workspace layout -----------------------------------------------------------
// [ tmp elements | bins D*P*256 u64 | plan | hist D*256 u64 | tickets D*P u32 | stats | look-back ]
//                                          ^--------- zeroed by one memset ---------------^
struct Layout {
    size_t tmp_off, bins_off, plan_off, hist_off, ticket_off, stats_off, lb_off, total;
    size_t zero_off, zero_bytes;
    int portions;
    size_t portion_elems;  // elements per portion (multiple of tile)
    size_t tiles_total;    // tiles over all portions (one pass)
};

Layout make_layout(size_t n, size_t elem_bytes, int digits, int tile) {
    Layout L{};
    const size_t max_portion = ((size_t)LB_VALUE_MASK / tile) * tile;  // < 2^30, whole tiles
    L.portion_elems = max_portion;
    L.portions = (int)((n + max_portion - 1) / max_portion);
    if (L.portions < 1) L.portions = 1;
    L.tiles_total = 0;
    for (int q = 0; q < L.portions; ++q) {
        const size_t pn = (q + 1 < L.portions) ? max_portion : n - (size_t)q * max_portion;
        L.tiles_total += (pn + tile - 1) / tile;
    }
    size_t off = 0;
    L.tmp_off = off;    off = align_up(off + n * elem_bytes, 256);
    L.bins_off = off;   off = align_up(off + (size_t)digits * L.portions * RADIX * 8, 256);
    L.plan_off = off;   off = align_up(off + sizeof(Plan), 256);
    L.hist_off = off;   off = align_up(off + (size_t)digits * RADIX * 8, 256);
    L.ticket_off = off; off = align_up(off + (size_t)digits * L.portions * 4, 256);
    L.stats_off = off;  off += 256;
    L.lb_off = off;     off = align_up(off + (size_t)digits * L.tiles_total * RADIX * 4, 256);
    L.total = off;
    L.zero_off = L.hist_off;
    L.zero_bytes = L.total - L.hist_off;
    return L;
}

struct PhaseTimer {  // optional CUDA-event timing of each phase (b200sort_keys_device_timed)
    std::vector<cudaEvent_t> ev;
    cudaStream_t st;
    bool on;
    explicit PhaseTimer(bool enable, cudaStream_t s) : st(s), on(enable) {}
    void mark() {
        if (!on) return;
        cudaEvent_t e;
        cudaEventCreate(&e);
        cudaEventRecord(e, st);
        ev.push_back(e);
    }
    ~PhaseTimer() { for (auto e : ev) cudaEventDestroy(e); }
};

// ---- the sort: launch sequence ---------------------------------------------------
template <typename Tr>
int sort_device_impl(typename Tr::elem_t* d_elems, void* d_tmp, size_t tmp_bytes, size_t n,
                     cudaStream_t st, unsigned flags, typename Tr::Params prm,
                     float* out_ms, int out_len, typename Tr::elem_t* alt_out = nullptr,
                     unsigned long long* d_hist_copy = nullptr, int hist_copy_words = 0) {
    using E = typename Tr::elem_t;
    constexpr int EB = (int)sizeof(E);
    constexpr int D = Tr::DIGITS;
    if (n == 0) return B200SORT_OK;
    if (!d_elems || !d_tmp) return fail(B200SORT_ERR_INVALID_ARG, "null device pointer");
    if (((uintptr_t)d_elems % alignof(E)) != 0) return fail(B200SORT_ERR_INVALID_ARG, "keys not aligned to the element size");
    if (((uintptr_t)d_tmp % 256) != 0) return fail(B200SORT_ERR_INVALID_ARG, "workspace not 256-byte aligned");

    const int variant = current_variant<EB>();
    const VariantDesc vd = Variants<EB>::v[variant];
    const int tile = vd.threads * vd.items;
    const Layout L = make_layout(n, EB, D, tile);
    if (tmp_bytes < L.total) return fail(B200SORT_ERR_INVALID_ARG, "workspace smaller than b200sort_temp_bytes()");

    unsigned char* ws = static_cast<unsigned char*>(d_tmp);
    // alt_out: single-pass callers (the multi-GPU partition) name the destination
    // themselves; the result stays there and nothing is copied back
    E* tmp = alt_out ? alt_out : reinterpret_cast<E*>(ws + L.tmp_off);
    unsigned long long* bins = reinterpret_cast<unsigned long long*>(ws + L.bins_off);
    Plan* plan = reinterpret_cast<Plan*>(ws + L.plan_off);
    unsigned long long* hist = reinterpret_cast<unsigned long long*>(ws + L.hist_off);
    uint32_t* tickets = reinterpret_cast<uint32_t*>(ws + L.ticket_off);
    uint32_t* lookback = reinterpret_cast<uint32_t*>(ws + L.lb_off);
    uint32_t* stats = reinterpret_cast<uint32_t*>(ws + L.stats_off);

    PhaseTimer timer(out_ms != nullptr, st);
    timer.mark();
    CU(cudaMemsetAsync(ws + L.zero_off, 0, L.zero_bytes, st));

    // K1: one read, all digit histograms
    {
        int dev = 0, sms = 148;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        constexpr int HT = 512;
        constexpr size_t HS = HistCfg<Tr>::SMEM;
        auto hk = hist_all_digits_kernel<Tr, HT>;
        static std::atomic<unsigned> configured{0};  // bit per device
        if (HS > 48 * 1024 && !(configured.load(std::memory_order_acquire) & (1u << (dev & 31)))) {
            CU(cudaFuncSetAttribute(hk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)HS));
            configured.fetch_or(1u << (dev & 31), std::memory_order_release);
        }
        const size_t per_sm = std::max<size_t>(1, std::min<size_t>(4, (220 * 1024) / HS));
        const size_t vecs = (n * EB + 15) / 16;
        size_t blocks = (vecs + (size_t)HT * 4 - 1) / ((size_t)HT * 4);
        const size_t cap = (size_t)sms * per_sm;
        if (blocks > cap) blocks = cap;
        if (blocks < 1) blocks = 1;
        hk<<<(unsigned)blocks, HT, HS, st>>>(d_elems, n, hist, prm);
        g_launches.fetch_add(1, std::memory_order_relaxed);
        CU(cudaGetLastError());
    }
    timer.mark();
    if (d_hist_copy)
        CU(cudaMemcpyAsync(d_hist_copy, hist, (size_t)hist_copy_words * 8, cudaMemcpyDeviceToDevice, st));
    // K2: bucket bases + pass plan
    scan_plan_kernel<<<1, RADIX, 0, st>>>(hist, bins, plan, (unsigned long long)n, D, L.portions,
                                          (flags & B200SORT_FLAG_NO_SKIP) ? 1 : 0);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    CU(cudaGetLastError());
    timer.mark();

    // K3: one launch per digit per portion
    size_t lb_cursor = 0;
    for (int p = 0; p < D; ++p) {
        for (int q = 0; q < L.portions; ++q) {
            const size_t off = (size_t)q * L.portion_elems;
            const size_t pn = (q + 1 < L.portions) ? L.portion_elems : n - off;
            PassArgs<Tr> a;
            a.keys = d_elems;
            a.tmp = tmp;
            a.plan = plan;
            a.pass = p;
            a.portion_off = off;
            a.portion_n = (uint32_t)pn;
            a.num_tiles = (uint32_t)((pn + tile - 1) / tile);
            a.bins_in = bins + ((size_t)p * L.portions + q) * RADIX;
            a.bins_out = (q + 1 < L.portions) ? bins + ((size_t)p * L.portions + q + 1) * RADIX : nullptr;
            a.lookback = lookback + lb_cursor * RADIX;
            a.ticket = tickets + (size_t)p * L.portions + q;
            a.stats = stats;
            a.prm = prm;
            lb_cursor += a.num_tiles;
            CU((launch_pass_idx<Tr, EB, 0>(variant, a, st)));
        }
        timer.mark();
    }
    // result in tmp after an odd number of executed passes (mod.rs:170-174)
    if (!alt_out) {
        const size_t bytes = n * EB;
        size_t blocks = (bytes / 16 + 255) / 256;
        if (blocks > 148 * 8) blocks = 148 * 8;
        if (blocks < 1) blocks = 1;
        copy_back_kernel<E><<<(unsigned)blocks, 256, 0, st>>>(d_elems, tmp, n, plan);
        g_launches.fetch_add(1, std::memory_order_relaxed);
        CU(cudaGetLastError());
    }
    timer.mark();

    if (out_ms) {
        if (out_len < D + 4) return fail(B200SORT_ERR_INVALID_ARG, "out_len < digits + 4");
        CU(cudaStreamSynchronize(st));
        // marks: 0 start | 1 after memset+K1 | 2 after K2 | 3..2+D after each pass | 3+D after copy-back
        for (int i = 0; i < D + 3; ++i) cudaEventElapsedTime(&out_ms[i], timer.ev[i], timer.ev[i + 1]);
        cudaEventElapsedTime(&out_ms[D + 3], timer.ev.front(), timer.ev.back());
    }
    return B200SORT_OK;
}

template <typename Tr>
size_t temp_bytes_impl(size_t n) {
    constexpr int EB = (int)sizeof(typename Tr::elem_t);
    // the maximum over all variants, so a sweep never outgrows the workspace
    size_t m = 0;
    for (int i = 0; i < num_variants<EB>(); ++i) {
        const size_t t = make_layout(n ? n : 1, EB, Tr::DIGITS, Variants<EB>::v[i].threads * Variants<EB>::v[i].items).total;
        if (t > m) m = t;
    }
    return m;
}

}  // namespace

// ---- kind dispatch ---------------------------------------------------------------
// One KeyTraits instantiation per built-in RadixDigits impl (radix_digits.rs:7-124).
#define B200_FOR_KIND(kind, FN, ...)                                                                  \
    switch (kind) {                                                                                   \
    case B200SORT_U8:   return FN<KeyTraits<uint8_t, XF_NONE>>(__VA_ARGS__);                          \
    case B200SORT_U16:  return FN<KeyTraits<uint16_t, XF_NONE>>(__VA_ARGS__);                         \
    case B200SORT_U32:  return FN<KeyTraits<uint32_t, XF_NONE>>(__VA_ARGS__);                         \
    case B200SORT_U64:  return FN<KeyTraits<unsigned long long, XF_NONE>>(__VA_ARGS__);               \
    case B200SORT_U128: return FN<KeyTraits<u128, XF_NONE>>(__VA_ARGS__);                             \
    case B200SORT_I8:   return FN<KeyTraits<uint8_t, XF_SIGNED>>(__VA_ARGS__);                        \
    case B200SORT_I16:  return FN<KeyTraits<uint16_t, XF_SIGNED>>(__VA_ARGS__);                       \
    case B200SORT_I32:  return FN<KeyTraits<uint32_t, XF_SIGNED>>(__VA_ARGS__);                       \
    case B200SORT_I64:  return FN<KeyTraits<unsigned long long, XF_SIGNED>>(__VA_ARGS__);             \
    case B200SORT_I128: return FN<KeyTraits<u128, XF_SIGNED>>(__VA_ARGS__);                           \
    case B200SORT_F32:  return FN<KeyTraits<uint32_t, XF_FLOAT>>(__VA_ARGS__);                        \
    case B200SORT_F64:  return FN<KeyTraits<unsigned long long, XF_FLOAT>>(__VA_ARGS__);              \
    default: break;                                                                                   \
    }

namespace {

template <typename Tr>
int keys_device_t(void* d_keys, void* d_tmp, size_t tmp_bytes, size_t n, cudaStream_t st,
                  unsigned flags, float* out_ms, int out_len) {
    return sort_device_impl<Tr>(static_cast<typename Tr::elem_t*>(d_keys), d_tmp, tmp_bytes, n, st,
                                flags, typename Tr::Params{}, out_ms, out_len);
}
template <typename Tr>
size_t keys_temp_t(size_t n) { return temp_bytes_impl<Tr>(n); }

int keys_device_dispatch(void* d_keys, void* d_tmp, size_t tmp_bytes, size_t n, int kind,
                         cudaStream_t st, unsigned flags, float* out_ms, int out_len) {
    B200_FOR_KIND(kind, keys_device_t, d_keys, d_tmp, tmp_bytes, n, st, flags, out_ms, out_len)
    return fail(B200SORT_ERR_UNSUPPORTED_KIND, "unknown key kind " + std::to_string(kind));
}

// ---- records: (K, V) tuples, key = one field (radix_digits.rs:126-136) ------------
int kind_xf(int kind) {
    switch (kind) {
    case B200SORT_I8: case B200SORT_I16: case B200SORT_I32: case B200SORT_I64: case B200SORT_I128: return XF_SIGNED;
    case B200SORT_F32: case B200SORT_F64: return XF_FLOAT;
    default: return XF_NONE;
    }
}

template <typename R, typename K>
int records_device_rk(void* d_recs, void* d_tmp, size_t tmp_bytes, size_t n, int key_shift, int xf,
                      cudaStream_t st) {
    switch (xf) {
    case XF_NONE: {
        using Tr = RecordTraits<R, K, XF_NONE>;
        return sort_device_impl<Tr>(static_cast<R*>(d_recs), d_tmp, tmp_bytes, n, st, 0, typename Tr::Params{key_shift}, nullptr, 0);
    }
    case XF_SIGNED: {
        using Tr = RecordTraits<R, K, XF_SIGNED>;
        return sort_device_impl<Tr>(static_cast<R*>(d_recs), d_tmp, tmp_bytes, n, st, 0, typename Tr::Params{key_shift}, nullptr, 0);
    }
    default: {
        using Tr = RecordTraits<R, K, XF_FLOAT>;
        return sort_device_impl<Tr>(static_cast<R*>(d_recs), d_tmp, tmp_bytes, n, st, 0, typename Tr::Params{key_shift}, nullptr, 0);
    }
    }
}

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

// ---- multi-GPU partition (K7): one onesweep pass keyed by destination rank ---------
template <typename Tr>
int partition_t(const void* d_in, void* d_out, void* d_tmp, size_t tmp_bytes, size_t n,
                const void* splitters, int nsplit, unsigned long long* d_counts, cudaStream_t st) {
    using U = typename Tr::elem_t;
    using PT = PartitionTraits<U, Tr::XF>;
    typename PT::Params prm{};
    prm.nsplit = nsplit;
    const U* raw = static_cast<const U*>(splitters);
    for (int j = 0; j < nsplit; ++j) prm.spl[j] = to_ordered<U, Tr::XF>(raw[j]);
    for (int j = 1; j < nsplit; ++j)
        if (ordered_less<U>(prm.spl[j], prm.spl[j - 1]))
            return fail(B200SORT_ERR_INVALID_ARG, "splitters are not ascending");
    if (n == 0) {
        CU(cudaMemsetAsync(d_counts, 0, (size_t)(nsplit + 1) * 8, st));
        return B200SORT_OK;
    }
    return sort_device_impl<PT>(static_cast<U*>(const_cast<void*>(d_in)), d_tmp, tmp_bytes, n, st,
                                B200SORT_FLAG_NO_SKIP, prm, nullptr, 0, static_cast<U*>(d_out), d_counts, nsplit + 1);
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

// copy in, sort, copy out; `sort` is called with (device elems, workspace, workspace bytes, stream)
template <typename SortFn>
int host_roundtrip(void* host, size_t bytes, size_t ws_bytes, SortFn sort) {
    const size_t data_bytes = align_up(bytes, 256);
    void* block = nullptr;
    int rc = pool_acquire(data_bytes + ws_bytes, &block);
    if (rc) return rc;
    StreamGuard sg;
    cudaError_t e = cudaStreamCreateWithFlags(&sg.st, cudaStreamNonBlocking);
    if (e != cudaSuccess) { pool_release(block); return cuda_fail(e, "cudaStreamCreate"); }
    unsigned char* d = static_cast<unsigned char*>(block);
    e = cudaMemcpyAsync(d, host, bytes, cudaMemcpyHostToDevice, sg.st);
    if (e == cudaSuccess) {
        rc = sort(d, d + data_bytes, ws_bytes, sg.st);
        if (rc == B200SORT_OK) e = cudaMemcpyAsync(host, d, bytes, cudaMemcpyDeviceToHost, sg.st);
    }
    const cudaError_t e2 = cudaStreamSynchronize(sg.st);
    pool_release(block);
    if (rc) return rc;
    if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpyAsync");
    if (e2 != cudaSuccess) return cuda_fail(e2, "cudaStreamSynchronize");
    return B200SORT_OK;
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
    B200_FOR_KIND(kind, keys_temp_t, n)
    return 0;
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

int b200sort_keys_host(void* keys, size_t n, int kind) {
    const size_t kb = b200sort_key_bytes(kind);
    if (kb == 0) return fail(B200SORT_ERR_UNSUPPORTED_KIND, "unknown key kind " + std::to_string(kind));
    if (n == 0) return B200SORT_OK;
    if (!keys) return fail(B200SORT_ERR_INVALID_ARG, "null host pointer");
    const size_t ws = b200sort_temp_bytes(n, kind);
    return host_roundtrip(keys, n * kb, ws, [&](void* d, void* w, size_t wb, cudaStream_t st) {
        return keys_device_dispatch(d, w, wb, n, kind, st, 0, nullptr, 0);
    });
}

size_t b200sort_records_temp_bytes(size_t n, size_t rec_bytes, int kind) {
    (void)kind;
    if (rec_bytes == 8) return temp_bytes_impl<RecordTraits<unsigned long long, uint32_t, XF_NONE>>(n);
    if (rec_bytes == 16) return temp_bytes_impl<RecordTraits<u128, unsigned long long, XF_NONE>>(n);
    return 0;
}

int b200sort_records_device(void* d_recs, void* d_tmp, size_t tmp_bytes, size_t n, size_t rec_bytes,
                            size_t key_off, int kind, void* stream) {
    const int rc = check_record_shape(rec_bytes, key_off, kind);
    if (rc) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (rec_bytes == 8)
        return records_device_rk<unsigned long long, uint32_t>(d_recs, d_tmp, tmp_bytes, n, (int)key_off * 8, kind_xf(kind), st);
    return records_device_rk<u128, unsigned long long>(d_recs, d_tmp, tmp_bytes, n, (int)key_off * 8, kind_xf(kind), st);
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
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    unsigned long long* counts = reinterpret_cast<unsigned long long*>(d_counts);
    B200_FOR_KIND(kind, partition_t, d_in, d_out, d_tmp, tmp_bytes, n, splitters, nsplit, counts, st)
    return fail(B200SORT_ERR_UNSUPPORTED_KIND, "unknown key kind " + std::to_string(kind));
}

unsigned long long b200sort_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int b200sort_set_variant(int key_bytes, int variant) {
    if (key_bytes != 1 && key_bytes != 2 && key_bytes != 4 && key_bytes != 8 && key_bytes != 16)
        return fail(B200SORT_ERR_INVALID_ARG, "key_bytes must be 1, 2, 4, 8 or 16");
    int nv = 0;
    switch (key_bytes) {
    case 1: nv = num_variants<1>(); break;
    case 2: nv = num_variants<2>(); break;
    case 4: nv = num_variants<4>(); break;
    case 8: nv = num_variants<8>(); break;
    default: nv = num_variants<16>(); break;
    }
    if (variant < -1 || variant >= nv) return fail(B200SORT_ERR_INVALID_ARG, "no such variant");
    g_variant[key_bytes].store(variant < 0 ? 0 : variant, std::memory_order_relaxed);
    return B200SORT_OK;
}

}  // extern "C"
