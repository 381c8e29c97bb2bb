// sort_impl.cuh -- host side of one sort: workspace layout and the launch
// sequence K1 -> K2 -> K3 x D -> copy-back, templated on the element traits.
//
// The library is compiled as many small translation units so that the build
// parallelises (radix_sort_b200/_build.py):
//   inst_front.cu    -DB200_FAM=f -DB200_W=w          sort_front<Tr>: K1/K2/copy-back
//                                                      + the launch sequence
//   inst_variant.cu  -DB200_FAM=f -DB200_W=w -DB200_V=v  one K3 instantiation
// and b200sort_api.cu dispatches the C ABI onto them.
#pragma once
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <string>
#include <vector>

#include "../../include/b200sort.h"
#include "narrow_kernels.cuh"
#include "sort_kernels.cuh"
#include "stream_kernel.cuh"

namespace b200 {

// ---- shared by all translation units (defined in b200sort_api.cu) ----------------
int fail(int code, const std::string& msg);
int cuda_fail(cudaError_t e, const char* what);
void count_launch();
int variant_override(int family, int elem_bytes);  // -1 = library default
bool force_rerank();                               // test hook (b200sort_debug_force_rerank)
bool small_sort_enabled();                         // test hook (b200sort_debug_small_sort)
bool narrow_fill_enabled();                        // test hook (b200sort_debug_narrow_fill)

#define B200_CU(call)                                               \
    do {                                                            \
        cudaError_t e_ = (call);                                    \
        if (e_ != cudaSuccess) return ::b200::cuda_fail(e_, #call); \
    } while (0)

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---- element families -------------------------------------------------------------
enum { FAM_KEYS = 0, FAM_RECORDS = 1, FAM_PARTITION = 2 };

template <int FAM, int W> struct FamilyTraits;
template <int W> struct FamilyTraits<FAM_KEYS, W> { using type = KeyTraits<typename UIntOf<W>::type>; };
template <int W> struct FamilyTraits<FAM_PARTITION, W> { using type = PartitionTraits<typename UIntOf<W>::type>; };
template <> struct FamilyTraits<FAM_RECORDS, 8> { using type = RecordTraits<unsigned long long, uint32_t>; };
template <> struct FamilyTraits<FAM_RECORDS, 16> { using type = RecordTraits<u128, unsigned long long>; };

// ---- K3 variants ------------------------------------------------------------------
// kernel: 0 = onesweep_pass_kernel (one tile per CTA)
//         3 = onesweep_stream_kernel (resident CTAs, next tile's keys prefetched into the dead
//             key registers; stream_kernel.cuh)
//         (1 and 2 were the TMA-ring kernels, removed: see sort_kernels.cuh)
// Index 0 of a table is the default (the fastest measured on B200, profiles/);
// The other entries exist
// for on-GPU sweeps (b200sort_set_variant).  The counts are repeated as macros
// because the dispatcher and the build script need them in the preprocessor.
struct VariantDesc { int kernel, threads, items, min_blocks, lbb; };  // lbb = predecessors read per look-back round (0 = LB_BATCH); bits 8+ = timing-ablation mask

template <int FAM, int W> struct Variants;
#define B200_NV_KEYS_1 2
template <> struct Variants<FAM_KEYS, 1> { static constexpr VariantDesc v[] = {{0, 512, 32, 2, 0x2000}, {3, 512, 32, 2, 0}}; };
#define B200_NV_KEYS_2 2
template <> struct Variants<FAM_KEYS, 2> { static constexpr VariantDesc v[] = {{0, 512, 24, 2, 0x2000}, {3, 512, 24, 2, 0}}; };
// (lbb bits 8+ = the ABL / FEAT mask of sort_kernels.cuh: FEAT_STG = 32 -> 0x2000, FEAT_NODOM = 64 -> 0x4000)
// Round-2 sweeps (profiles/r02_sweep_v30_v32.README.txt, r02_sweep_v33_stg_1B.jsonl): st.global write-out +2.9 % per middle
// pass on u32 (default there and for the narrower widths), but on u64 the last pass loses more than the middle ones win
// (1 B keys: 37.21 ms against 36.97) -> plain stores stay the u64 default;
// plain-atomic scatter (FEAT_NODOM), 1024-thread tiles with one CTA per SM (-9 %), 32- and 40-register shapes with
// 64 / 48 resident warps (-19 % / -1..8 %) all lose and are not shipped (1024x26x1, 1024x16x1, 1024x8x2, 768x10x2 ...).
#define B200_NV_KEYS_4 6
template <> struct Variants<FAM_KEYS, 4> { static constexpr VariantDesc v[] = {{0, 512, 26, 2, 0x2000}, {3, 512, 22, 2, 0}, {0, 512, 22, 2, 0}, {0, 512, 16, 2, 0},
    {0, 512, 26, 2, 0x6000}, {0, 512, 26, 2, 0}}; };
#define B200_NV_KEYS_8 5
template <> struct Variants<FAM_KEYS, 8> { static constexpr VariantDesc v[] = {{0, 512, 16, 2, 0}, {3, 512, 16, 2, 0}, {0, 256, 20, 3, 0},
    {0, 512, 16, 2, 0x6000}, {0, 512, 16, 2, 0x2000}}; };
#define B200_NV_KEYS_16 2
template <> struct Variants<FAM_KEYS, 16> { static constexpr VariantDesc v[] = {{0, 256, 12, 2, 0x2000}, {3, 256, 12, 2, 0}}; };
#define B200_NV_RECORDS_8 2
template <> struct Variants<FAM_RECORDS, 8> { static constexpr VariantDesc v[] = {{0, 512, 12, 2, 0x2000}, {3, 512, 12, 2, 0}}; };
#define B200_NV_RECORDS_16 2
template <> struct Variants<FAM_RECORDS, 16> { static constexpr VariantDesc v[] = {{0, 256, 12, 2, 0x2000}, {3, 256, 12, 2, 0}}; };
// (the partition is a single, non-stable pass 0: its write-out never verifies, so FEAT_STG would be masked out anyway)
#define B200_NV_PARTITION_1 1
template <> struct Variants<FAM_PARTITION, 1> { static constexpr VariantDesc v[] = {{0, 512, 32, 2, 0}}; };
#define B200_NV_PARTITION_2 1
template <> struct Variants<FAM_PARTITION, 2> { static constexpr VariantDesc v[] = {{0, 512, 24, 2, 0}}; };
#define B200_NV_PARTITION_4 1
template <> struct Variants<FAM_PARTITION, 4> { static constexpr VariantDesc v[] = {{0, 512, 16, 2, 0}}; };
#define B200_NV_PARTITION_8 1
template <> struct Variants<FAM_PARTITION, 8> { static constexpr VariantDesc v[] = {{0, 512, 12, 2, 0}}; };
#define B200_NV_PARTITION_16 1
template <> struct Variants<FAM_PARTITION, 16> { static constexpr VariantDesc v[] = {{0, 256, 12, 2, 0}}; };

template <int FAM, int W> constexpr int num_variants() { return (int)(sizeof(Variants<FAM, W>::v) / sizeof(VariantDesc)); }

// One K3 launch of variant V; member defined in variant_def.cuh and instantiated
// by inst_variant.cu (one translation unit per variant).
template <int FAM, int W, int V>
struct VariantEntry {
    static cudaError_t launch(const PassArgs<typename FamilyTraits<FAM, W>::type>& a, cudaStream_t st);
};

template <typename Tr, int KERNEL, int THREADS, int ITEMS, int MINB, int LBB_ABL>
struct PassLauncher {
    using SM = PassSmem<Tr, THREADS, ITEMS>;
    static_assert(KERNEL == 0 || KERNEL == 3, "kernel ids: 0 one tile per CTA, 3 streaming");
    static constexpr int BLOCK = THREADS;
    static constexpr int LBB = (LBB_ABL & 255) > 0 ? (LBB_ABL & 255) : LB_BATCH;  // bits 8+: timing-ablation mask
    static constexpr size_t SMEM = SM::TOTAL;
    static auto kernel() {
        if constexpr (KERNEL == 3) return onesweep_stream_kernel<Tr, THREADS, ITEMS, MINB, LBB, (LBB_ABL >> 8)>;
        else return onesweep_pass_kernel<Tr, THREADS, ITEMS, MINB, LBB, (LBB_ABL >> 8)>;
    }
    // per-device one-time setup: opt in to the dynamic shared memory, read the occupancy
    static cudaError_t prepare(int* resident) {
        static std::atomic<int> cached[32];  // resident CTAs + 1; 0 = not prepared on that device
        int dev = 0;
        cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return e;
        int c = cached[dev & 31].load(std::memory_order_acquire);
        if (c == 0) {
            auto k = kernel();
            e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM);
            if (e != cudaSuccess) return e;
            int per_sm = 0, sms = 0;
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, BLOCK, SMEM);
            if (e != cudaSuccess) return e;
            e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            if (e != cudaSuccess) return e;
            if (per_sm < 1) return cudaErrorLaunchOutOfResources;
            c = per_sm * sms + 1;
            cached[dev & 31].store(c, std::memory_order_release);
        }
        *resident = c - 1;
        return cudaSuccess;
    }
    static cudaError_t launch(const PassArgs<Tr>& a, cudaStream_t st) {
        int resident = 0;
        cudaError_t e = prepare(&resident);
        if (e != cudaSuccess) return e;
        unsigned grid = a.num_tiles;
        if (KERNEL == 3 && grid > (unsigned)resident) grid = (unsigned)resident;
        kernel()<<<grid, BLOCK, SMEM, st>>>(a);
        count_launch();
        return cudaGetLastError();
    }
};

// ---- workspace layout -----------------------------------------------------------
// [ tmp elements | bins D*P*256 u64 | plan | hist D*256 u64 | tickets D*P u32 | stats | look-back ]
// (the partition family always writes to a caller-named destination: no tmp region)
//                                          ^--------- zeroed by one memset ---------------^
struct Layout {
    size_t tmp_off, bins_off, plan_off, hist_off, ticket_off, stats_off, lb_off, narrow_off, narrow_partial_off, total;
    size_t zero_off, zero_bytes;
    int portions;
    size_t portion_elems;  // elements per portion (multiple of tile)
    size_t tiles_total;    // tiles over all portions (one pass)
};

// narrow_bins: 256 / 65536 for key-only 1- / 2-byte keys (value histogram + prefix of the counting
// path, narrow_kernels.cuh: NV u64 bins + NV + 1 u64 offsets, after everything else), else 0
inline Layout make_layout(size_t n, size_t elem_bytes, int digits, int tile, bool with_tmp = true, size_t narrow_bins = 0) {
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
    L.tmp_off = off;    off = align_up(off + (with_tmp ? n * elem_bytes : 0), 256);
    L.bins_off = off;   off = align_up(off + (size_t)digits * L.portions * RADIX * 8, 256);
    L.plan_off = off;   off = align_up(off + sizeof(Plan), 256);
    L.hist_off = off;   off = align_up(off + (size_t)digits * RADIX * 8, 256);
    L.ticket_off = off; off = align_up(off + (size_t)digits * L.portions * 4, 256);
    L.stats_off = off;  off += 256;
    L.lb_off = off;     off = align_up(off + (size_t)(with_tmp ? 2 : 1) * L.tiles_total * RADIX * 4, 256);  // two halves, swapped per executed pass
    L.zero_off = L.hist_off;
    L.zero_bytes = off - L.hist_off;
    L.narrow_off = off; off = align_up(off + (narrow_bins ? (2 * narrow_bins + 1) * 8 : 0), 256);
    // u16: per-slab partial histograms (narrow_kernels.cuh), sized for the largest slab count any device gets for n
    L.narrow_partial_off = off;
    if (narrow_bins == 65536) off = align_up(off + (size_t)narrow_slabs16(n, 2 * NARROW_MAX_SLABS) * narrow_bins * 4, 256);
    L.total = off;
    return L;
}

struct PhaseTimer {  // optional CUDA-event timing of each phase (b200sort_keys_device_timed)
    std::vector<cudaEvent_t> ev;
    cudaStream_t st;
    bool on;
    PhaseTimer(bool enable, cudaStream_t s) : st(s), on(enable) {}
    void mark() {
        if (!on) return;
        cudaEvent_t e;
        cudaEventCreate(&e);
        cudaEventRecord(e, st);
        ev.push_back(e);
    }
    ~PhaseTimer() { for (auto e : ev) cudaEventDestroy(e); }
};

// One sort request, type-erased so that b200sort_api.cu needs no kernel templates.
struct SortCall {
    void* elems = nullptr;        // device: keys / records (in place) or partition input
    void* ws = nullptr;           // device workspace, 256-byte aligned
    size_t ws_bytes = 0;
    size_t n = 0;
    cudaStream_t st = nullptr;
    unsigned flags = 0;
    int xf = XF_NONE;             // key transform (radix_digits.rs: unsigned / signed / float)
    int key_shift = 0;            // records: bit offset of the key inside the record
    float* out_ms = nullptr;      // per-phase timing (synchronises) or null
    int out_len = 0;
    void* alt_out = nullptr;      // partition: destination buffer (single pass, no copy-back)
    unsigned long long* hist_copy = nullptr;  // partition: per-destination counts (device)
    int hist_copy_words = 0;
    const void* splitters = nullptr;          // partition: nsplit raw keys (host)
    int nsplit = 0;
    // partition, split in two calls around the caller's count exchange (fused multi-GPU path):
    //   PHASE_COUNT   memset + K1 + K2: leaves histogram, plan and look-back state in ws
    //   PHASE_SCATTER K3 only, on the state PHASE_COUNT left in the same ws; bucket r goes to
    //                 peer_bases[r] (peer memory) starting at element peer_offsets[r]
    //   PHASE_READ_STATS: no kernel; synchronises and copies the workspace's counter of re-ranked tiles to *out_stats
    //   PHASE_CLAIM   K3 only, no counts at all: every tile claims room in its buckets' regions with an atomic on
    //                 claim_cursors[r] (device addresses, possibly on peers) -- PartitionTraits::Params::claim
    enum { PHASE_ALL = 0, PHASE_COUNT = 1, PHASE_SCATTER = 2, PHASE_CLAIM = 3, PHASE_READ_STATS = 6 };
    const unsigned long long* claim_cursors = nullptr;  // host array, nsplit + 1 device addresses of u64 cursors
    const unsigned long long* claim_caps = nullptr;     // host array, nsplit + 1 region capacities (elements)
    void* claim_flags = nullptr;                        // device u32: bit r set when bucket r did not fit
    unsigned* out_stats = nullptr;
    int phase = PHASE_ALL;
    unsigned bucket_mask = 0xFFFFFFFFu;              // PHASE_SCATTER: buckets this launch stores (bit r = bucket r)
    bool rezero = false;                             // PHASE_SCATTER run again on the same state: zero tickets + look-back first
    void* const* peer_bases = nullptr;               // host array, nsplit + 1 device pointers
    const unsigned long long* peer_offsets = nullptr;  // device array, nsplit + 1 element offsets
};

// entry points of the per-(family, width) front translation units
template <int FAM, int W> int sort_front(const SortCall& c);
template <int FAM, int W> size_t temp_bytes_front(size_t n);

}  // namespace b200
