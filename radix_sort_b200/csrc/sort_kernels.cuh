// sort_kernels.cuh -- the sm_100a kernels of the LSD radix sort.
//
//   K1 hist_all_digits_kernel  one read of the keys -> all D x 256 histograms
//                              (replaces phase A, reference mod.rs:90-109; the
//                              all-digits-up-front idea is the reference's own
//                              radix_sort1, mod.rs:217-239)
//   K2 scan_plan_kernel        exclusive scan of every histogram -> bucket bases,
//                              plus the pass plan (replaces phase B, mod.rs:110-120)
//   K3 onesweep_pass_kernel    one digit pass: stage tile -> rank -> decoupled
//                              look-back -> coalesced write-out (replaces phase C,
//                              the 96-element buffered scatter, mod.rs:121-168)
//   K4 = the key transform, fused into K1 and into the first/last executed K3
//   copy_back_kernel           odd number of executed passes (mod.rs:170-174)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "key_traits.cuh"

namespace b200 {

constexpr int RADIX = 256;          // 8-bit digits, radix_digits.rs:4 (get_digit -> u8)
constexpr int MAX_DIGITS = 16;      // u128
constexpr unsigned FULL_MASK = 0xFFFFFFFFu;

// Look-back status word: {flag:2, value:30}.  A portion holds < 2^30 keys so an
// inclusive prefix always fits.
constexpr uint32_t LB_FLAG_PARTIAL = 1u << 30;
constexpr uint32_t LB_FLAG_INCLUSIVE = 1u << 31;
constexpr uint32_t LB_VALUE_MASK = (1u << 30) - 1u;

struct PassPlan {
    uint32_t skip;        // digit is constant over all keys: pass is the identity
    uint32_t src_is_tmp;  // ping-pong side this pass reads (mod.rs:85-89)
    uint32_t xf_in;       // first executed pass: map raw keys -> ordered keys on load
    uint32_t xf_out;      // last executed pass: map ordered keys -> raw keys on store
};
struct Plan {
    PassPlan pass[MAX_DIGITS];
    uint32_t result_in_tmp;  // odd number of executed passes
    uint32_t n_exec;
};

__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// read-only element load (LDG.E.*.CONSTANT); 128-bit elements go as one LDG.128
template <typename E>
__device__ __forceinline__ E ld_elem(const E* p) { return __ldg(p); }
template <>
__device__ __forceinline__ u128 ld_elem<u128>(const u128* p) {
    const ulonglong2 v = __ldg(reinterpret_cast<const ulonglong2*>(p));
    u128 r;
    r.lo = v.x;
    r.hi = v.y;
    return r;
}

// store into global memory through the global state space (STG), not the generic one (ST):
// the address of a run comes out of integer arithmetic, which nvcc otherwise stores through
// the generic window check.  Peer-mapped receive buffers (K7) are global addresses too.
// No "memory" clobber: the write-out keeps its shared-memory loads of the next rows in flight
// across these stores (with the clobber pass 0 of a u64 sort was 4 % slower); every reader of
// the stored data is behind a barrier or a kernel boundary.  What this relies on: write_out is the LAST
// user of the destination buffer inside the kernel -- nothing after it loads from or stores to the addresses
// written here with ordinary C++ accesses, which the compiler would be free to move across these asm stores.
// (Keep it that way, or give the last store of each thread a "memory" clobber.)
__device__ __forceinline__ void st_global(unsigned long long addr, unsigned char v) { asm volatile("st.global.u8 [%0], %1;" ::"l"(addr), "r"((unsigned)v)); }
__device__ __forceinline__ void st_global(unsigned long long addr, unsigned short v) { asm volatile("st.global.u16 [%0], %1;" ::"l"(addr), "h"(v)); }
__device__ __forceinline__ void st_global(unsigned long long addr, unsigned int v) { asm volatile("st.global.u32 [%0], %1;" ::"l"(addr), "r"(v)); }
__device__ __forceinline__ void st_global(unsigned long long addr, unsigned long long v) { asm volatile("st.global.u64 [%0], %1;" ::"l"(addr), "l"(v)); }
__device__ __forceinline__ void st_global(unsigned long long addr, u128 v) { asm volatile("st.global.v2.u64 [%0], {%1, %2};" ::"l"(addr), "l"(v.lo), "l"(v.hi)); }

// ---------------------------------------------------------------------------
// K1: all-digit histogram.  16-byte coalesced loads (LDG.128), four in flight
// per thread; one privatised shared-memory histogram set per CTA, bumped with
// no-return shared atomics.  Measured on B200 (profiles/r01_*): a no-return ATOMS
// is not slowed by same-address lanes (the unit merges them), so skewed inputs
// need no software warp aggregation -- the ballot/branch version of this kernel
// spent 68 instructions per 32 keys for 21 % of HBM peak.  What does cost is bank
// conflicts between DIFFERENT digits, so every bin has R copies, the copy chosen
// by lane % R, with R as large as 128 KB of shared memory allows: R = 32 (keys of
// <= 4 digits) puts lane l on bank l for every digit -- one wavefront per atomic,
// no conflicts at all; with R = 16 (8-byte keys) lanes l and l+16 collide whenever
// their digits have the same parity -- two wavefronts per atomic, the LSU data pipe
// at 95 % of its peak, 3.9 TB/s (profiles/r01_ncu_k1_final_u64_2p27.txt); R = 8 for
// 16-byte keys.  Tried for 8-byte keys and dropped: two 16-bit counters per word
// (R = 32 in 16 KB per digit, flushed every 253 rounds) -- conflict-free, but the atomic
// then adds 1 or 1 << 16, and a generic ATOMS.ADD costs 3.3 cycles per warp instruction
// where the +1 form (ATOMS.POPC.INC) costs 1-2: 2.8 TB/s (r01_sweep_v23).
// ---------------------------------------------------------------------------
template <typename Tr>
struct HistCfg {
    static constexpr int THREADS = 1024;
    static constexpr int R = Tr::DIGITS <= 4 ? 32 : Tr::DIGITS <= 8 ? 16 : 8;
    static constexpr size_t SMEM = (size_t)Tr::DIGITS * RADIX * R * sizeof(uint32_t);
    static constexpr int VEC = 16 / (int)sizeof(typename Tr::elem_t);
};

template <typename Tr, int THREADS>
__global__ void __launch_bounds__(THREADS)
hist_all_digits_kernel(const typename Tr::elem_t* __restrict__ in, size_t n,
                       unsigned long long* __restrict__ hist, typename Tr::Params prm) {
    using E = typename Tr::elem_t;
    constexpr int D = Tr::DIGITS;
    constexpr int R = HistCfg<Tr>::R;
    constexpr int VEC = HistCfg<Tr>::VEC;
    extern __shared__ __align__(16) uint32_t hist_sh[];
    const int tid = threadIdx.x;
    for (int i = tid; i < D * RADIX * R; i += THREADS) hist_sh[i] = 0;
    __syncthreads();
    uint32_t* mine = hist_sh + (tid & (R - 1));

    auto count = [&](E raw) {
        const E e = Tr::xform_in(raw, prm);
#pragma unroll
        for (int p = 0; p < D; ++p) atomicAdd(&mine[(p * RADIX + Tr::digit(e, p, prm)) * R], 1u);
    };
    auto count_vec = [&](uint4 v) {
        union { uint4 v; E e[VEC]; } u;
        u.v = v;
#pragma unroll
        for (int j = 0; j < VEC; ++j) count(u.e[j]);
    };

    // peel to 16-byte alignment, then whole vectors, then the tail
    const uintptr_t addr = (uintptr_t)in;
    size_t head = ((16 - (addr & 15)) & 15) / sizeof(E);
    if (head > n) head = n;
    const uint4* vec = reinterpret_cast<const uint4*>(in + head);
    const size_t nvec = (n - head) / VEC;
    const size_t tail = head + nvec * VEC;

    // A round = four vectors per thread, `stride` apart; `rounds` = the rounds in which EVERY
    // thread of this CTA has all four (a CTA-uniform trip count: the loop control stays in
    // uniform registers; u32 K1 0.90 -> 0.82 ms at 1 B keys against a per-thread count); at
    // most four vectors per thread are left for the strided loop below.  Software pipeline:
    // the four vectors of the next round are requested before the current four are counted,
    // so HBM and the atomic unit work at the same time.
    const size_t stride = (size_t)gridDim.x * THREADS;
    size_t i = (size_t)blockIdx.x * THREADS + tid;
    const size_t cta_last = (size_t)blockIdx.x * THREADS + (THREADS - 1) + 3 * stride;  // the CTA's last vector of round 0
    const size_t rounds = nvec > cta_last ? (nvec - 1 - cta_last) / (4 * stride) + 1 : 0;
    if (rounds) {
        uint4 a = __ldg(vec + i), b = __ldg(vec + i + stride), c = __ldg(vec + i + 2 * stride), d = __ldg(vec + i + 3 * stride);
        for (size_t r = 1;; ++r) {
            i += 4 * stride;
            const bool more = r < rounds;
            uint4 na, nb, nc, nd;
            if (more) {
                na = __ldg(vec + i); nb = __ldg(vec + i + stride); nc = __ldg(vec + i + 2 * stride); nd = __ldg(vec + i + 3 * stride);
            }
            count_vec(a); count_vec(b); count_vec(c); count_vec(d);
            if (!more) break;
            a = na; b = nb; c = nc; d = nd;
        }
    }
    for (; i < nvec; i += stride) count_vec(__ldg(vec + i));
    if (blockIdx.x == 0) {
        for (size_t j = tid; j < head; j += THREADS) count(ld_elem(in + j));
        for (size_t j = tail + tid; j < n; j += THREADS) count(ld_elem(in + j));
    }
    __syncthreads();
    for (int k = tid; k < D * RADIX; k += THREADS) {
        uint32_t c = 0;
#pragma unroll
        for (int r = 0; r < R; ++r) c += hist_sh[k * R + ((r + tid) & (R - 1))];  // rotated: lanes read distinct banks
        if (c) atomicAdd(&hist[k], (unsigned long long)c);
    }
}

// ---------------------------------------------------------------------------
// K2: one CTA of 256 threads.  For every digit: exclusive scan of its 256-bin
// histogram -> global bucket bases of portion 0 (64-bit); detection of constant
// digits; then thread 0 lays out the pass plan (ping-pong side, where the key
// transform is applied, where the result ends up).
// bins layout: [digit][portion][256].
// ---------------------------------------------------------------------------
static __global__ void __launch_bounds__(RADIX)
scan_plan_kernel(const unsigned long long* __restrict__ hist, unsigned long long* __restrict__ bins,
                 Plan* __restrict__ plan, unsigned long long n, int digits, int portions, int no_skip) {
    __shared__ unsigned long long warp_tot[RADIX / 32];
    __shared__ uint32_t s_skip[MAX_DIGITS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int p = 0; p < digits; ++p) {
        const unsigned long long c = hist[p * RADIX + tid];
        unsigned long long x = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long y = __shfl_up_sync(FULL_MASK, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_tot[warp] = x;
        const int is_const = __syncthreads_or(c == n);
        unsigned long long base = 0;
        for (int w = 0; w < warp; ++w) base += warp_tot[w];
        bins[((size_t)p * portions) * RADIX + tid] = base + x - c;
        if (tid == 0) s_skip[p] = (is_const && !no_skip) ? 1u : 0u;
        __syncthreads();
    }
    if (tid == 0) {
        int first = -1, last = -1, n_exec = 0;
        for (int p = 0; p < digits; ++p)
            if (!s_skip[p]) { if (first < 0) first = p; last = p; }
        for (int p = 0; p < MAX_DIGITS; ++p) {
            PassPlan pp;
            pp.skip = (p < digits) ? s_skip[p] : 1u;
            pp.src_is_tmp = (uint32_t)(n_exec & 1);
            pp.xf_in = (p == first);
            pp.xf_out = (p == last);
            plan->pass[p] = pp;
            if (!pp.skip) ++n_exec;
        }
        plan->n_exec = (uint32_t)n_exec;
        plan->result_in_tmp = (uint32_t)(n_exec & 1);
    }
}

// ---------------------------------------------------------------------------
// K3: one digit pass over one portion (< 2^30 elements), onesweep style.
//
// Ranking.  Measured on B200 (profiles/r01_microbench_*.txt, r01 ncu captures):
// MATCH.ANY on 8-bit digits costs 41-47 cycles per warp instruction per SM (ADU
// pipe, ~1 cycle per distinct value) and capped the first version of this kernel
// at 0.46 keys/cycle/SM; a shared-memory atomicAdd costs ~3.5 wavefronts.  So:
//   1. count:   every key bumps its warp's private counter of its digit (no return);
//   2. bases:   thread d turns column d of the warp counters into each warp's first
//               slot for digit d inside the digit-sorted tile, and publishes the
//               tile's count of d for the look-back;
//   3. scatter: a second atomicAdd, WITH return, hands every key its final slot in
//               the tile in shared memory -- no per-key rank registers at all.
// Step 3 is a correct LSD pass iff the lanes of one warp instruction that hit the
// same counter are served in ascending lane order: true in 3e10 checked
// lane-rounds on B200 (profiles/r01_microbench_atomics_order.txt) but not an
// architectural promise, so the kernel does not rely on it:
//   * key-only sorts need less than stability -- equal keys are bit-identical, so
//     a pass is correct iff the digit-sorted tile is non-decreasing on digits
//     0..pass.  The write-out loop checks exactly that (one neighbour read per
//     key); a tile that fails is re-ranked by rerank_tile() and rewritten, and
//     counted in stats[0].
//   * records (NEEDS_STABLE) need stability proper, and equal keys with different
//     payloads cannot be told apart by comparing keys.  So the scatter also drops one
//     word per tile slot, (digit << 16) | source slot of the record, into sh.ent; the
//     tile is stably sorted by digit iff those words rise strictly along it, which the
//     write-out checks (two 1-wavefront loads per warp instruction).  A failing tile
//     takes the same slow path.  (Ranking every record tile with ballots -- 8 ballots
//     give the lanes holding the same digit, lane-ordered by construction -- was the
//     first version: (u32,u32) 39.7 -> 46.9 Grecords/s without it.)
//   * rerank_tile(), the slow path, ranks with those ballots.  Tests force it
//     (b200sort_debug_force_rerank) and read the counter (b200sort_debug_reranked_tiles).
//
// Look-back.  With C tiles in flight and a tile started every g = T_tile / C, a
// walk that pays one L2 round trip l per predecessor settles at a depth of about
// l / g predecessors -- i.e. l^2 C / T_tile of pure latency per tile, which was as
// long as the rest of the tile (r01 ncu: the spin loop held 45 % of the stall
// samples).  So the walk reads LB_BATCH predecessors per round trip (independent
// relaxed loads), and the first batch is requested before the scatter so that it
// has landed by the time the scatter is done.
//
// Two kernels share this tile code (rank_and_scatter):
//   onesweep_pass_kernel    one tile per CTA, keys loaded straight into registers.
//   onesweep_stream_kernel  (stream_kernel.cuh) resident CTAs loop over tickets and
//                           keep the next tile's loads in flight.
// Two TMA-ring kernels (cp.async.bulk + mbarrier, look-back in dedicated warps,
// deferred write-out) were built and measured 10-40 % slower on every width
// (profiles/r01_sweep_v5*, v13*, v16*); they were removed after commit 823d302 --
// the polling warps and the second tile buffer cost more than the overlap gained.
// ---------------------------------------------------------------------------
template <typename Tr>
struct PassArgs {
    typename Tr::elem_t* keys;           // caller's buffer
    typename Tr::elem_t* tmp;            // ping-pong buffer (mod.rs:71-83)
    const Plan* plan;
    int pass;                            // digit index (mod.rs:84)
    unsigned long long portion_off;      // first element of this portion
    uint32_t portion_n;                  // elements in this portion
    uint32_t num_tiles;
    uint32_t grid;                       // CTAs to launch (streaming kernel: resident CTAs)
    const unsigned long long* bins_in;   // [256] global base of each bucket for this portion
    unsigned long long* bins_out;        // [256] base for the next portion (or null)
    uint32_t* lookback;                  // [2][num_tiles][256] status words, zeroed.  An executed pass uses half
                                         // plan.src_is_tmp (the parity of the passes executed before it) and
                                         // zeroes the other half for the next executed pass
    size_t lookback_half;                // words between the halves; 0 = one half only (single-pass partition)
    uint32_t* ticket;                    // tile-id dispenser, zeroed
    uint32_t* stats;                     // [0] tiles that took the re-rank fallback
    uint32_t force_rerank;               // test hook: every tile takes the re-rank fallback (b200sort_debug_force_rerank)
    uint32_t opaque_zero;                // always 0, but ptxas cannot know: see pass_s in rank_and_scatter
    typename Tr::Params prm;
};

constexpr int LB_BATCH = 8;

// Bits of the ABL template parameter of the K3 kernels (VariantDesc::lbb >> 8).  Bits 0..4 are timing
// ablations (results wrong on purpose); the bits from 5 up switch code paths that give the same result.
constexpr int FEAT_STG = 32;     // write-out through st.global (STG) instead of generic stores (ST)
constexpr int FEAT_NODOM = 64;   // scatter: plain atomics unless the tile or the warp has a dominant digit

template <typename Tr, int THREADS, int ITEMS>
struct PassSmem {
    static constexpr int WARPS = THREADS / 32;
    static constexpr int TILE = THREADS * ITEMS;
    static constexpr size_t TILE_BYTES = (size_t)TILE * sizeof(typename Tr::elem_t);
    static constexpr size_t HIST_BYTES = (size_t)WARPS * RADIX * sizeof(uint32_t);
    static constexpr size_t GPTR_BYTES = (size_t)RADIX * sizeof(unsigned long long);
    static constexpr size_t EXCL_BYTES = (size_t)RADIX * sizeof(uint32_t);
    static constexpr size_t CNT_BYTES = (size_t)RADIX * sizeof(uint32_t);
    // records: one word {digit:16, source slot:16} per tile slot, the stability check of the write-out
    static constexpr size_t ENT_BYTES = Tr::NEEDS_STABLE ? (size_t)TILE * sizeof(uint32_t) : 0;
    static_assert(!Tr::NEEDS_STABLE || TILE <= 65536, "source slots are kept in 16 bits");
    static constexpr size_t AUX_BYTES = HIST_BYTES + GPTR_BYTES + EXCL_BYTES + CNT_BYTES + 64 + 16 + ENT_BYTES;
    static constexpr size_t TOTAL = TILE_BYTES + AUX_BYTES;
};

// views into the dynamic shared memory of one CTA: [tile buffers | aux]
template <typename Tr, int THREADS, int ITEMS>
struct PassShared {
    using SM = PassSmem<Tr, THREADS, ITEMS>;
    typename Tr::elem_t* tile;      // [TILE]   digit-sorted tile
    uint32_t* warp_hist;            // [WARPS][256] per-warp digit counters
    unsigned long long* gptr;       // [256] byte address in dst of slot 0 of the tile, per digit
    uint32_t* excl;                 // [256] keys of the tile with a smaller digit
    uint32_t* cnt;                  // [256] keys of the tile with this digit
    uint32_t* misc;                 // [0..1] tickets, [2..9] warp totals of the digit scan
    uint32_t* ent;                  // [TILE] records only: (digit << 16) | source slot of the record in this tile slot
    __device__ __forceinline__ PassShared() {}
    __device__ __forceinline__ PassShared(unsigned char* raw, size_t tile_bytes_total) {
        tile = reinterpret_cast<typename Tr::elem_t*>(raw);
        unsigned char* aux = raw + tile_bytes_total;
        warp_hist = reinterpret_cast<uint32_t*>(aux);
        gptr = reinterpret_cast<unsigned long long*>(aux + SM::HIST_BYTES);
        excl = reinterpret_cast<uint32_t*>(aux + SM::HIST_BYTES + SM::GPTR_BYTES);
        cnt = excl + RADIX;
        misc = cnt + RADIX;
        ent = misc + 20;  // misc holds 64 + 16 bytes
    }
};

// Counter of digit d inside a 256-counter histogram.  Counter index = bank, so
// digits that are multiples of 16 or 32 (stride-2^k sequences, floats made from
// small integers) would pile onto one or two banks; folding the top three bits
// into the low five spreads them (a bijection; r01 sweep v14: ascending u32,
// pass 0, 1.54 -> 1.04 ms).
__device__ __forceinline__ unsigned hidx(unsigned d) { return d ^ (d >> 5); }

// lane-ordered rank of one item inside its warp: 8 ballots -> lanes with my digit
template <bool FULL_TILE>
__device__ __forceinline__ uint32_t ballot_rank(unsigned d, bool valid, uint32_t* my_hist, int lane, unsigned lane_lt) {
    unsigned m = FULL_TILE ? FULL_MASK : __ballot_sync(FULL_MASK, valid);
#pragma unroll
    for (int b = 0; b < 8; ++b) {
        const bool bit = (d >> b) & 1u;
        const unsigned bal = __ballot_sync(FULL_MASK, bit);
        m &= bit ? bal : ~bal;
    }
    const int leader = __ffs(m) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(&my_hist[hidx(d)], (uint32_t)__popc(m));
    base = __shfl_sync(FULL_MASK, base, leader & 31);
    return base + __popc(m & lane_lt);
}

// Thread d owns digit d: the tile's count of d, and the per-warp first slots.
template <int WARPS>
__device__ __forceinline__ uint32_t column_total(const uint32_t* warp_hist, int d) {
    uint32_t c = 0;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) c += warp_hist[w * RADIX + hidx(d)];
    return c;
}
template <int WARPS>
__device__ __forceinline__ void column_bases(uint32_t* warp_hist, int d, uint32_t start) {
    uint32_t run = start;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) {
        const uint32_t t = warp_hist[w * RADIX + hidx(d)];
        warp_hist[w * RADIX + hidx(d)] = run;
        run += t;
    }
}

// Make the compiler forget what it knows about a key (no instruction is emitted).
// Without this it keeps every key's counter ADDRESS from the count phase alive for
// the scatter phase -- ITEMS extra registers, which at the 64-register cap end up in
// local memory (STL/LDL in the hot loop: more LSU traffic on an LSU-bound kernel).
// Recomputing shift + mask + add per key is cheaper.
__device__ __forceinline__ void launder(unsigned char& v) { unsigned t = v; asm volatile("" : "+r"(t)); v = (unsigned char)t; }
__device__ __forceinline__ void launder(unsigned short& v) { asm volatile("" : "+h"(v)); }
__device__ __forceinline__ void launder(unsigned int& v) { asm volatile("" : "+r"(v)); }
__device__ __forceinline__ void launder(unsigned long long& v) { asm volatile("" : "+l"(v)); }
__device__ __forceinline__ void launder(u128& v) { asm volatile("" : "+l"(v.lo), "+l"(v.hi)); }

// Decoupled look-back for one digit, LB_BATCH predecessors per round trip.
// p points at the status word of tile `pos` (the nearest predecessor) for this
// digit; tiles below 0 read as {INCLUSIVE, 0}.
template <int B>
struct LookBack {
    uint32_t v[B];
    __device__ __forceinline__ void request(const uint32_t* p, int pos) {
#pragma unroll
        for (int j = 0; j < B; ++j)
            v[j] = (pos - j >= 0) ? ld_relaxed_u32(p - (size_t)j * RADIX) : LB_FLAG_INCLUSIVE;
    }
    // sum of the counts of all preceding tiles.  Branch-free inside a batch: the walk is
    // a serial chain run by 8 of the CTA's 16 warps while the others wait at the
    // barrier (r01 ncu, u64: a quarter of the tile's wall time, most of it the
    // per-word branches, not the round trips), so a batch costs a handful of
    // selects per word and ONE branch.  A batch that holds an unpublished word in
    // front of its first INCLUSIVE word is asked for again as a whole.
    __device__ __forceinline__ uint32_t resolve(const uint32_t* p, int pos) {
        uint32_t sum = 0;
        for (;;) {
            uint32_t part = 0;
            bool open = true, ok = true;
#pragma unroll
            for (int j = 0; j < B; ++j) {
                const uint32_t x = v[j];
                ok &= !open || x >= LB_FLAG_PARTIAL;       // a word of the open range not yet published
                part += open ? (x & LB_VALUE_MASK) : 0u;
                open &= (x & LB_FLAG_INCLUSIVE) == 0;
            }
            if (ok) {
                sum += part;
                if (!open) return sum;
                p -= (size_t)B * RADIX;
                pos -= B;
            }
            request(p, pos);
        }
    }
};

// value of lane `src` of the warp, for every element width
__device__ __forceinline__ unsigned char shfl_elem(unsigned char v, int src) { return (unsigned char)__shfl_sync(FULL_MASK, (unsigned)v, src); }
__device__ __forceinline__ unsigned short shfl_elem(unsigned short v, int src) { return (unsigned short)__shfl_sync(FULL_MASK, (unsigned)v, src); }
__device__ __forceinline__ unsigned int shfl_elem(unsigned int v, int src) { return __shfl_sync(FULL_MASK, v, src); }
__device__ __forceinline__ unsigned long long shfl_elem(unsigned long long v, int src) { return __shfl_sync(FULL_MASK, v, src); }
__device__ __forceinline__ u128 shfl_elem(u128 v, int src) {
    u128 r;
    r.lo = __shfl_sync(FULL_MASK, v.lo, src);
    r.hi = __shfl_sync(FULL_MASK, v.hi, src);
    return r;
}

// Coalesced run write-out of the digit-sorted tile; VERIFY adds the order check
// (every key against its successor in the tile).  Two slot-to-thread mappings,
// chosen by element width from what B200 measured (profiles/r01_sweep_v19*):
//   * elements of 8 bytes and more: warp w owns slots [w*32*ITEMS, (w+1)*32*ITEMS),
//     row j of it = 32 consecutive slots, one per lane.  The successor of a key
//     sits in the next lane's register, so it comes by shuffle instead of a second
//     shared-memory load (2 wavefronts for a 64-bit LDS): lane 0 contributes its
//     NEXT row's key (what lane 31 needs), the other lanes their own, and lane l
//     reads from lane l+1 mod 32.  u64 passes 1.164 -> 1.144 ms at 2^28 keys.
//   * narrower elements: thread t handles slots t, t+THREADS, ...; the successor is
//     one more 1-wavefront load -- a shuffle is dearer than that (u32 0.756 vs 0.770).
// (32-bit element offsets instead of the 64-bit addresses in sh.gptr were tried:
// one wavefront less per warp instruction, but no gain -- v19.)
template <typename E>
__device__ __forceinline__ void st_elem(unsigned long long addr, E v, bool stg) {
    if (stg) st_global(addr, v);
    else *reinterpret_cast<E*>(addr) = v;
}

template <typename Tr, int THREADS, int ITEMS, bool FULL, bool VERIFY, int FEAT = 0>
__device__ __forceinline__ bool write_out(const PassShared<Tr, THREADS, ITEMS>& sh, const PassArgs<Tr>& a,
                                          uint32_t xf_out, uint32_t tile_n, int tid) {
    using E = typename Tr::elem_t;
    bool bad = false;
    constexpr bool STG = (FEAT & FEAT_STG) != 0;
    const uint32_t n_valid = FULL ? (uint32_t)(THREADS * ITEMS) : tile_n;
    if constexpr (Tr::FEW_BUCKETS) {
        // long runs, expensive digit: copy run by run, no digit per key
        const int buckets = a.prm.nsplit + 1;
        for (int b = 0; b < buckets; ++b) {
            if (!((a.prm.mask >> b) & 1u)) continue;
            const uint32_t lo = sh.excl[b], hi = lo + sh.cnt[b];
            const unsigned long long base = sh.gptr[b];
            if (base == 0) continue;  // claim mode: the region is full (the caller falls back)
            for (uint32_t idx = lo + tid; idx < hi; idx += THREADS)
                st_elem<E>(base + (unsigned long long)idx * sizeof(E), sh.tile[idx], STG);
        }
    } else if constexpr (sizeof(E) >= 8) {
        const int lane = tid & 31;
        const uint32_t wb = (uint32_t)(tid >> 5) * 32u * ITEMS + (uint32_t)lane;
        // rolling window of two rows: the compiler hoists as many of the (independent)
        // loads as it has registers for -- all of them in the one-tile kernel, few in the
        // streaming kernel, whose key registers already hold the next tile
        E cur{};
        if (FULL || wb < n_valid) cur = sh.tile[wb];
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            const uint32_t idx = wb + j * 32;
            E below{};  // the key one row down; past the last row: the next warp's first key (lane 0 only)
            if (j + 1 < ITEMS) {
                if (FULL || idx + 32 < n_valid) below = sh.tile[idx + 32];
            } else if (VERIFY && !Tr::NEEDS_STABLE) {
                if (lane == 0 && idx + 32 < n_valid) below = sh.tile[idx + 32];
            }
            E key = cur;
            if constexpr (VERIFY && Tr::NEEDS_STABLE) {
                // records: {digit, source slot} must rise strictly along the tile (stable by digit)
                if (idx + 1 < n_valid) bad |= sh.ent[idx] >= sh.ent[idx + 1];
            } else if constexpr (VERIFY) {
                const E nxt = shfl_elem(lane == 0 ? below : key, (lane + 1) & 31);
                if (idx + 1 < n_valid) bad |= Tr::prefix_less(nxt, key, a.pass, a.prm);
            }
            if (FULL || idx < n_valid) {
                const unsigned d = Tr::digit(key, a.pass, a.prm);
                if (xf_out) key = Tr::xform_out(key, a.prm);
                st_elem<E>(sh.gptr[d] + (unsigned long long)idx * sizeof(E), key, STG);
            }
            cur = below;
        }
    } else {
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            const uint32_t idx = (uint32_t)j * THREADS + tid;
            if (FULL || idx < n_valid) {
                E key = sh.tile[idx];
                if constexpr (VERIFY && Tr::NEEDS_STABLE) {
                    if (idx + 1 < n_valid) bad |= sh.ent[idx] >= sh.ent[idx + 1];
                } else if constexpr (VERIFY) {
                    if (idx + 1 < n_valid) bad |= Tr::prefix_less(sh.tile[idx + 1], key, a.pass, a.prm);
                }
                const unsigned d = Tr::digit(key, a.pass, a.prm);
                if (xf_out) key = Tr::xform_out(key, a.prm);
                st_elem<E>(sh.gptr[d] + (unsigned long long)idx * sizeof(E), key, STG);
            }
        }
    }
    return bad;
}

// Slow path, out of line: rebuild the tile with ballot ranks (lane-ordered by
// construction) from the keys still in global memory, and rewrite its slots.
// Entered by a whole CTA; warp_hist may hold anything; excl/gptr are valid.
template <typename Tr, int THREADS, int ITEMS>
__device__ __noinline__ void rerank_tile(const PassShared<Tr, THREADS, ITEMS> sh,
                                         const PassArgs<Tr>* ap, const typename Tr::elem_t* src_tile,
                                         uint32_t tile_n, uint32_t xf_in, uint32_t xf_out) {
    using E = typename Tr::elem_t;
    constexpr int WARPS = THREADS / 32;
    const PassArgs<Tr>& a = *ap;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned lane_lt = (1u << lane) - 1u;
    const uint32_t wbase = (uint32_t)warp * 32u * ITEMS + (uint32_t)lane;
    uint32_t* my_hist = sh.warp_hist + warp * RADIX;
    __syncthreads();
    for (int i = tid; i < WARPS * RADIX; i += THREADS) sh.warp_hist[i] = 0;
    __syncthreads();
    E e[ITEMS];
    uint32_t rank[ITEMS];
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        const bool valid = wbase + i * 32 < tile_n;
        if (valid) {
            e[i] = ld_elem(src_tile + wbase + i * 32);
            if (xf_in) e[i] = Tr::xform_in(e[i], a.prm);
        }
        rank[i] = ballot_rank<false>(Tr::digit(e[i], a.pass, a.prm), valid, my_hist, lane, lane_lt);
    }
    __syncthreads();
    if (tid < RADIX) column_bases<WARPS>(sh.warp_hist, tid, sh.excl[tid]);
    __syncthreads();
#pragma unroll
    for (int i = 0; i < ITEMS; ++i)
        if (wbase + i * 32 < tile_n) sh.tile[rank[i] + my_hist[hidx(Tr::digit(e[i], a.pass, a.prm))]] = e[i];
    __syncthreads();
    write_out<Tr, THREADS, ITEMS, false, false>(sh, a, xf_out, tile_n, tid);
    __syncthreads();
}

// One tile whose keys are already in registers (warp-striped: item i of lane l =
// element i*32+l of the warp's contiguous segment, so slot order == memory order).
// FULL (all THREADS*ITEMS slots valid) is known at compile time so the per-key
// loops carry no bounds predicates.  sh.tile may alias the buffer the keys were
// read from: the first barrier below orders those reads before the scatter.
// Returns true when the order check of the write-out failed (caller re-ranks).
struct NoHook { __device__ __forceinline__ void operator()() const {} };

// WORKERS threads rank the tile; the last 256 of them also run the look-back.
template <typename Tr, int WORKERS, int ITEMS, bool FULL, int LBB, int ABL = 0, typename Hook = NoHook,
          typename Hook2 = NoHook>
__device__ __forceinline__ bool rank_and_scatter(const PassShared<Tr, WORKERS, ITEMS>& sh, const PassArgs<Tr>& a,
                                                 const PassPlan& pp, typename Tr::elem_t (&e)[ITEMS],
                                                 typename Tr::elem_t* __restrict__ dst, uint32_t tile, uint32_t tile_n,
                                                 Hook after_count = Hook{}, Hook2 after_scatter = Hook2{}) {
    using E = typename Tr::elem_t;
    constexpr int WARPS = WORKERS / 32;
    constexpr int LB0 = WORKERS - RADIX;  // in-line look-back: the last 256 threads
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int pass = a.pass;
    const uint32_t wbase = (uint32_t)warp * 32u * ITEMS + (uint32_t)lane;

    if (pp.xf_in) {
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) e[i] = Tr::xform_in(e[i], a.prm);
    }

    // ---- 1. count: the warp's counters become its digit histogram
    uint32_t* my_hist = sh.warp_hist + warp * RADIX;
    unsigned code[Tr::FEW_BUCKETS ? (ITEMS + 7) / 8 : 1] = {};  // FEW_BUCKETS: 4-bit digit of every key
    if constexpr (Tr::FEW_BUCKETS) {
#pragma unroll
        for (int i = 0; i < ITEMS; ++i)
            if (FULL || wbase + i * 32 < tile_n) {
                const unsigned d = Tr::digit(e[i], pass, a.prm);
                code[i / 8] |= d << (4 * (i % 8));
                atomicAdd(&my_hist[hidx(d)], 1u);
            }
    } else {
#pragma unroll
        for (int i = 0; i < ITEMS; ++i)
            if (FULL || wbase + i * 32 < tile_n) atomicAdd(&my_hist[hidx(Tr::digit(e[i], pass, a.prm))], 1u);
    }
    __syncthreads();
    if constexpr (!Tr::FEW_BUCKETS) {
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) launder(e[i]);
    }
    after_count();  // streaming kernel: thread 0 parks the next ticket

    // ---- decoupled look-back in line, one digit per thread of the last 256 threads,
    // run AFTER the scatter.  (Walking before the scatter publishes INCLUSIVE words
    // sooner but meets predecessors that have not published yet: 5-30 % slower with the
    // branchy walk, profiles/r01_sweep_v12_early_lookback.jsonl; with the branch-free
    // walk it is a tie, profiles/r01_sweep_v25_early_lookback_again.jsonl.)
    uint32_t* lb = a.lookback + (pp.src_is_tmp ? a.lookback_half : 0) + (size_t)tile * RADIX;
    auto lookback_inline = [&]() {
        if (tid >= LB0) {
            const int d = tid - LB0;
            const uint32_t cnt_d = sh.cnt[d];
            const unsigned long long bin_d = a.bins_in[d];  // asked for before the walk: its round trip overlaps the walk's
            unsigned long long excl_tiles = 0;
            if constexpr (Tr::FEW_BUCKETS) {
                if (a.prm.claim) {
                    unsigned long long base = 0;
                    if (d <= a.prm.nsplit && cnt_d != 0 && ((a.prm.mask >> d) & 1u)) {
                        const unsigned long long old =
                            atomicAdd_system(reinterpret_cast<unsigned long long*>(a.prm.cursor[d]), (unsigned long long)cnt_d);
                        if (old + cnt_d <= a.prm.cap[d])
                            base = Tr::dst_base((unsigned)d, dst, a.prm) + old * sizeof(E) - (unsigned long long)sh.excl[d] * sizeof(E);
                        else
                            atomicOr(reinterpret_cast<unsigned*>(a.prm.flags), 1u << d);
                    }
                    sh.gptr[d] = base;
                    return;
                }
            }
            if (ABL & 1) {  // timing ablation: no polling, a plausible distinct address
                excl_tiles = min((unsigned long long)tile * (WORKERS * ITEMS / RADIX),
                                 (unsigned long long)a.portion_n - cnt_d - min((unsigned long long)a.portion_n - cnt_d, bin_d));
            } else if (tile > 0) {
                LookBack<LBB> look;
                look.request(lb - RADIX + d, (int)tile - 1);
                const uint32_t sum = look.resolve(lb - RADIX + d, (int)tile - 1);
                st_relaxed_u32(lb + d, LB_FLAG_INCLUSIVE | (sum + cnt_d));
                excl_tiles = sum;
            }
            const unsigned long long gbase = bin_d + excl_tiles;
            sh.gptr[d] = Tr::dst_base((unsigned)d, dst, a.prm) + (gbase - sh.excl[d]) * sizeof(E);
            if (a.bins_out != nullptr && tile == a.num_tiles - 1) a.bins_out[d] = gbase + cnt_d;
        }
    };

    // ---- 2. thread d owns digit d: tile count -> publish -> scan over digits ->
    // per-warp first slots inside the digit-sorted tile
    uint32_t count = 0;
    if (tid < RADIX) {
        count = column_total<WARPS>(sh.warp_hist, tid);
        bool publish = true;
        if constexpr (Tr::FEW_BUCKETS) publish = !a.prm.claim;  // claim mode: nobody looks back
        if (publish) st_relaxed_u32(lb + tid, (tile == 0 ? LB_FLAG_INCLUSIVE : LB_FLAG_PARTIAL) | count);
        // the same tile's words in the other half: nobody reads them before the next launch
        if (a.lookback_half) a.lookback[(pp.src_is_tmp ? 0 : a.lookback_half) + (size_t)tile * RADIX + tid] = 0u;
        sh.cnt[tid] = count;
        // a digit that holds more than half of the tile (at most one can) is ranked
        // without atomics in the scatter; the word is tagged with tile id and pass so
        // that a stale one (shared memory survives launches) is not taken for this
        // tile's -- and if it ever were, any digit is a correct choice, only slower
        if (count * 2 > tile_n) sh.misc[10] = ((tile + 1u) << 12) | ((uint32_t)(pass & 15) << 8) | (uint32_t)tid;
    }
    uint32_t incl = count;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(FULL_MASK, incl, o);
        if (lane >= o) incl += y;
    }
    // (bits 16+: how many of this warp's 32 digits occur in the tile at all)
    const unsigned occupied = __ballot_sync(FULL_MASK, count != 0);
    if (lane == 31 && warp < RADIX / 32) sh.misc[2 + warp] = incl | ((uint32_t)__popc(occupied) << 16);
    __syncthreads();
    if (tid < RADIX) {
        uint32_t base = 0;
#pragma unroll
        for (int w = 0; w < RADIX / 32; ++w) base += (w < warp) ? (sh.misc[2 + w] & 0xFFFFu) : 0u;
        const uint32_t excl = base + incl - count;  // keys of the tile with a smaller digit
        sh.excl[tid] = excl;
        column_bases<WARPS>(sh.warp_hist, tid, excl);
    }
    __syncthreads();

    // ---- 3. scatter: the returned counter value IS the key's slot in the tile.
    // A returning shared atomic serialises lanes that hit the SAME counter (25 cycles
    // per warp instruction when all 32 do, 14 when two counters share them, 4 from
    // 16 counters up -- profiles/r01_microbench_atomics_skew.txt), which is what
    // clustered, sorted, few-unique or heavy-headed inputs produce.  So the keys of
    // the tile's dominant digit (if one holds more than half of the tile) take no
    // atomic at all: one ballot gives their lane-ordered rank inside the instruction,
    // a running count in a register the rest -- lane-ordered by construction.  The
    // other keys are few and take the atomic.  The partition kernel (2..16 buckets,
    // registers to spare) groups lanes with match_any when a tile has <= 3 buckets.
    // launder() stops nvcc from keeping every key's counter address alive from the count phase,
    // but ptxas still sees the same registers and value-numbers digit(e[i], pass) of both phases
    // alike (spills at the 64-register cap).  The scatter extracts its digits with a shift amount
    // ptxas cannot prove equal: pass ^ 0, the zero coming in as a kernel parameter.
    const int pass_s = pass ^ (int)a.opaque_zero;
    if constexpr (Tr::FEW_BUCKETS) {
        uint32_t distinct = 0;
        if constexpr (FULL) {
#pragma unroll
            for (int w = 0; w < RADIX / 32; ++w) distinct += sh.misc[2 + w] >> 16;
        }
        if (FULL && distinct <= 3) {
            const unsigned lane_lt = (1u << lane) - 1u;
#pragma unroll
            for (int i = 0; i < ITEMS; ++i) {
                const unsigned d = (code[Tr::FEW_BUCKETS ? i / 8 : 0] >> (4 * (i % 8))) & 15u;
                const unsigned m = __match_any_sync(FULL_MASK, d);
                const int leader = __ffs(m) - 1;
                uint32_t base = 0;
                if (lane == leader) base = atomicAdd(&my_hist[hidx(d)], (uint32_t)__popc(m));
                base = __shfl_sync(FULL_MASK, base, leader);
                sh.tile[base + __popc(m & lane_lt)] = e[i];
            }
        } else {
#pragma unroll
            for (int i = 0; i < ITEMS; ++i)
                if (FULL || wbase + i * 32 < tile_n) {
                    const unsigned d = (code[Tr::FEW_BUCKETS ? i / 8 : 0] >> (4 * (i % 8))) & 15u;
                    sh.tile[atomicAdd(&my_hist[hidx(d)], 1u)] = e[i];
                }
        }
    } else if constexpr ((ABL & 8) != 0) {  // timing ablation: identity slots
#pragma unroll
        for (int i = 0; i < ITEMS; ++i)
            if (FULL || wbase + i * 32 < tile_n) sh.tile[wbase + i * 32] = e[i];
    } else {
        const uint32_t domw = sh.misc[10];
        // no digit above half of the tile: the digit of the warp's first key stands in
        // (any choice is correct; on sorted input, whose tiles straddle two runs, it
        // is the warp's own dominant digit)
        const bool tile_dom = (domw >> 8) == ((((tile + 1u) << 4) | (uint32_t)(pass & 15)) & 0xFFFFFFu);
        const unsigned d0 = Tr::digit(e[0], pass_s, a.prm);
        const unsigned dom = tile_dom ? (domw & 0xFFu) : __shfl_sync(FULL_MASK, d0, 0);
        // FEAT_NODOM: the ballot path costs every key ~10 instructions and a divergent
        // branch around its atomic (the loop cannot be software-pipelined), which keys
        // without a dominant digit pay for nothing.  Warp-uniform choice: ballot path
        // if the tile has a dominant digit or a quarter of the warp's first row agrees
        // with its first key (sorted / clustered input); plain atomics otherwise.
        bool use_dom = true;
        if constexpr ((ABL & FEAT_NODOM) != 0) use_dom = tile_dom || __popc(__ballot_sync(FULL_MASK, d0 == dom)) >= 8;
        if (use_dom) {
            uint32_t dom_slot = my_hist[hidx(dom)];  // next free slot of that digit in this warp's range
            const unsigned lane_lt = (1u << lane) - 1u;
#pragma unroll
            for (int i = 0; i < ITEMS; ++i) {
                const bool valid = FULL || wbase + i * 32 < tile_n;
                const unsigned d = Tr::digit(e[i], pass_s, a.prm);
                const bool is_dom = valid && d == dom;
                const unsigned m = __ballot_sync(FULL_MASK, is_dom);
                uint32_t slot = dom_slot + (uint32_t)__popc(m & lane_lt);
                dom_slot += (uint32_t)__popc(m);
                if (valid && !is_dom) slot = atomicAdd(&my_hist[hidx(d)], 1u);
                if (valid) {
                    sh.tile[slot] = e[i];
                    if constexpr (Tr::NEEDS_STABLE) sh.ent[slot] = (d << 16) | (wbase + (uint32_t)i * 32u);
                }
            }
        } else {
            // groups of SG keys: SG atomics in flight, then SG stores (the whole tile at once
            // would need ITEMS slot registers on top of the keys: spills at the 64-register cap)
            constexpr int SG = sizeof(E) >= 8 ? 4 : 6;
#pragma unroll
            for (int g = 0; g < ITEMS; g += SG) {
                uint32_t slot[SG];
#pragma unroll
                for (int j = 0; j < SG; ++j) {
                    const int i = g + j;
                    if (i < ITEMS && (FULL || wbase + i * 32 < tile_n))
                        slot[j] = atomicAdd(&my_hist[hidx(Tr::digit(e[i], pass_s, a.prm))], 1u);
                }
#pragma unroll
                for (int j = 0; j < SG; ++j) {
                    const int i = g + j;
                    if (i < ITEMS && (FULL || wbase + i * 32 < tile_n)) {
                        sh.tile[slot[j]] = e[i];
                        if constexpr (Tr::NEEDS_STABLE)
                            sh.ent[slot[j]] = (Tr::digit(e[i], pass_s, a.prm) << 16) | (wbase + (uint32_t)i * 32u);
                    }
                }
                asm volatile("" ::: "memory");
            }
        }
    }
    after_scatter();  // e[] is dead from here on: the streaming kernel re-fills it with the next tile

    lookback_inline();
    __syncthreads();

    // ---- coalesced run write-out; key-only passes check the order of the tile
    if (ABL & 4) return false;
    if ((!Tr::NEEDS_STABLE && pass == 0) || (ABL & 2)) {
        // (FEAT_STG only pays in the verifying write-out: pass 0 is 3-4 % slower with it -- r02 sweep v33)
        write_out<Tr, WORKERS, ITEMS, FULL, false, (ABL & ~FEAT_STG)>(sh, a, pp.xf_out, tile_n, tid);
        return false;
    }
    return write_out<Tr, WORKERS, ITEMS, FULL, true, ABL>(sh, a, pp.xf_out, tile_n, tid);
}

template <typename Tr, int THREADS, int ITEMS, int MIN_BLOCKS, int LBB = LB_BATCH, int ABL = 0>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS)
onesweep_pass_kernel(const __grid_constant__ PassArgs<Tr> a) {
    using E = typename Tr::elem_t;
    using SM = PassSmem<Tr, THREADS, ITEMS>;
    constexpr int WARPS = THREADS / 32;
    constexpr int TILE = THREADS * ITEMS;
    static_assert(THREADS >= RADIX && THREADS % 32 == 0, "one thread per digit needed");

    // the ticket and the pass plan are two independent round trips to L2: ask for both at once
    uint32_t ticket = 0;
    if (threadIdx.x == 0) ticket = atomicAdd(a.ticket, 1u);
    const PassPlan pp = a.plan->pass[a.pass];
    if (pp.skip) return;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    PassShared<Tr, THREADS, ITEMS> sh(smem_raw, SM::TILE_BYTES);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    const E* __restrict__ src = (pp.src_is_tmp ? a.tmp : a.keys) + a.portion_off;
    E* __restrict__ dst = pp.src_is_tmp ? a.keys : a.tmp;

    // tile id from a ticket, not blockIdx: every predecessor a tile may wait on
    // in the look-back has already started, whatever order CTAs are dispatched in.
    // (Requesting the keys of tile blockIdx.x before the ticket is known was tried:
    // tickets and block ids disagree often enough that the re-loads cost more than
    // the hidden round trip -- r01 sweep v7.)
    if (tid == 0) sh.misc[0] = ticket;
#pragma unroll
    for (int i = tid; i < WARPS * RADIX; i += THREADS) sh.warp_hist[i] = 0;
    __syncthreads();
    const uint32_t tile = sh.misc[0];
    const uint32_t tile_n = min((uint32_t)TILE, a.portion_n - tile * (uint32_t)TILE);
    const E* src_tile = src + (size_t)tile * TILE;
    const uint32_t wbase = (uint32_t)warp * 32u * ITEMS + (uint32_t)lane;
    E e[ITEMS];
    bool bad;
    if (tile_n == (uint32_t)TILE) {
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) e[i] = ld_elem(src_tile + wbase + i * 32);
        bad = rank_and_scatter<Tr, THREADS, ITEMS, true, LBB, ABL>(sh, a, pp, e, dst, tile, tile_n, NoHook{});
    } else {
#pragma unroll
        for (int i = 0; i < ITEMS; ++i)
            if (wbase + i * 32 < tile_n) e[i] = ld_elem(src_tile + wbase + i * 32);
        bad = rank_and_scatter<Tr, THREADS, ITEMS, false, LBB, ABL>(sh, a, pp, e, dst, tile, tile_n, NoHook{});
    }
    if ((ABL & 31) != 0) return;  // (bits 0..4 are timing ablations: results wrong on purpose)
    if (!Tr::NEEDS_STABLE && a.pass == 0 && !a.force_rerank) return;  // no order to check, and bad == false
    if (!__syncthreads_or(bad || a.force_rerank)) return;
    if (tid == 0) atomicAdd(a.stats, 1u);
    rerank_tile<Tr, THREADS, ITEMS>(sh, &a, src_tile, tile_n, pp.xf_in, pp.xf_out);
}

// ---------------------------------------------------------------------------
// Small sorts: n <= SMALL_THREADS * SMALL_ITEMS elements in ONE launch of one CTA.
// The multi-kernel sequence (memset, K1, K2, D x K3, copy-back) is launch-bound
// below ~1e5 keys (75 us for 1000 u32 keys, 125 us for u64: r01 small-n sweep); a
// drop-in for slice.radix_sort() is called on small slices too.  The keys stay in
// registers across the passes; every pass ranks them with the ballot ranker (stable
// by construction, so the same code serves records), scatters them into shared
// memory and reads them back in slot order.  A digit on which all keys agree is
// skipped, as K2 does for the big path.
// ---------------------------------------------------------------------------
constexpr int SMALL_THREADS = 512;
constexpr int SMALL_ITEMS = 8;

template <typename Tr>
struct SmallSmem {
    static constexpr int TILE = SMALL_THREADS * SMALL_ITEMS;
    static constexpr size_t TILE_BYTES = (size_t)TILE * sizeof(typename Tr::elem_t);
    static constexpr size_t TOTAL = TILE_BYTES + (size_t)(SMALL_THREADS / 32) * RADIX * sizeof(uint32_t) + 64;
};

template <typename Tr>
__global__ void __launch_bounds__(SMALL_THREADS, 1)
small_sort_kernel(typename Tr::elem_t* keys, uint32_t n, typename Tr::Params prm, int no_skip) {
    using E = typename Tr::elem_t;
    constexpr int ITEMS = SMALL_ITEMS;
    constexpr int WARPS = SMALL_THREADS / 32;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    E* tile = reinterpret_cast<E*>(smem_raw);
    uint32_t* warp_hist = reinterpret_cast<uint32_t*>(smem_raw + SmallSmem<Tr>::TILE_BYTES);
    uint32_t* misc = warp_hist + WARPS * RADIX;  // [0..7] warp totals of the digit scan
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned lane_lt = (1u << lane) - 1u;
    const uint32_t wbase = (uint32_t)warp * 32u * ITEMS + (uint32_t)lane;  // warp-striped: slot order == memory order
    uint32_t* my_hist = warp_hist + warp * RADIX;

    E e[ITEMS];
#pragma unroll
    for (int i = 0; i < ITEMS; ++i)
        if (wbase + i * 32 < n) e[i] = Tr::xform_in(keys[wbase + i * 32], prm);  // plain load: the kernel stores into the same buffer

    for (int pass = 0; pass < Tr::DIGITS; ++pass) {
#pragma unroll
        for (int i = lane; i < RADIX; i += 32) my_hist[i] = 0;
        __syncwarp();
        uint32_t rank[ITEMS];
#pragma unroll
        for (int i = 0; i < ITEMS; ++i)
            rank[i] = ballot_rank<false>(Tr::digit(e[i], pass, prm), wbase + i * 32 < n, my_hist, lane, lane_lt);
        __syncthreads();
        // thread d owns digit d: count over the warps, exclusive scan over the digits
        uint32_t count = 0, incl = 0;
        if (tid < RADIX) count = column_total<WARPS>(warp_hist, tid);
        incl = count;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(FULL_MASK, incl, o);
            if (lane >= o) incl += y;
        }
        if (lane == 31 && warp < RADIX / 32) misc[warp] = incl;
        const bool constant = __syncthreads_or(tid < RADIX && count == n) != 0;  // one digit holds every key
        if (tid < RADIX) {
            uint32_t base = 0;
#pragma unroll
            for (int w = 0; w < RADIX / 32; ++w) base += (w < warp) ? misc[w] : 0u;
            column_bases<WARPS>(warp_hist, tid, base + incl - count);
        }
        __syncthreads();
        if (constant && !no_skip) continue;  // the pass is the identity (CTA-uniform)
#pragma unroll
        for (int i = 0; i < ITEMS; ++i)
            if (wbase + i * 32 < n) tile[rank[i] + my_hist[hidx(Tr::digit(e[i], pass, prm))]] = e[i];
        __syncthreads();
#pragma unroll
        for (int i = 0; i < ITEMS; ++i)
            if (wbase + i * 32 < n) e[i] = tile[wbase + i * 32];
        // (the next scatter into `tile` is two barriers away)
    }
#pragma unroll
    for (int i = 0; i < ITEMS; ++i)
        if (wbase + i * 32 < n) keys[wbase + i * 32] = Tr::xform_out(e[i], prm);
}

// ---------------------------------------------------------------------------
// Odd number of executed passes: the result sits in tmp (mod.rs:170-174).
// Exits at once when the plan says the result is already in the caller's buffer.
// ---------------------------------------------------------------------------
template <typename E>
__global__ void __launch_bounds__(256)
copy_back_kernel(E* __restrict__ keys, const E* __restrict__ tmp, size_t n, const Plan* __restrict__ plan) {
    if (!plan->result_in_tmp) return;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const size_t t0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if ((((uintptr_t)keys | (uintptr_t)tmp) & 15) == 0) {
        constexpr size_t VEC = 16 / sizeof(E);
        const size_t nvec = n / VEC;
        const uint4* s = reinterpret_cast<const uint4*>(tmp);
        uint4* d = reinterpret_cast<uint4*>(keys);
        for (size_t i = t0; i < nvec; i += stride) d[i] = s[i];
        for (size_t i = nvec * VEC + t0; i < n; i += stride) keys[i] = tmp[i];
    } else {
        for (size_t i = t0; i < n; i += stride) keys[i] = tmp[i];
    }
}

}  // namespace b200
