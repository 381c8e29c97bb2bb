// ring_kernel.cuh -- K3, third form: persistent CTAs, three-slot TMA ring,
// write-out deferred by one tile, look-back in dedicated warps.
//
// Why.  ncu on the two earlier forms (profiles/r01_k3_*): a third of the stall
// samples sit in, or at the barrier behind, the decoupled look-back -- the tile
// cannot be written out before its global offsets are known, and those cost 2-3
// dependent round trips to L2.  Starting the look-back earlier does not help:
// the walk then meets predecessors that have not published yet.  So the
// dependence is removed from the critical path instead:
//
//   workers, iteration i        look-back warps
//   ------------------------    ------------------------------------------
//   wait TMA(tile i)
//   count -> publish partial i  (BAR_CNT)  walk back for tile i, publish
//   scan, scatter tile i        (BAR_EXCL) INCLUSIVE(i), gptr[i & 1]
//   write out tile i-1   <----- (BAR_GPTR, arrived one tile-time ago)
//   start TMA(tile i+2) into the slot tile i-1 just left
//
// The walk for tile i has until the workers finish ranking tile i+1, and the
// INCLUSIVE words still appear as early as they can, which keeps everybody's
// walk short.  Costs: a third ring slot, double-buffered cnt/excl/gptr.
#pragma once
#include "sort_kernels.cuh"

namespace b200 {

template <typename Tr, int WORKERS, int ITEMS>
struct RingSmem {
    static constexpr int WARPS = WORKERS / 32;
    static constexpr int TILE = WORKERS * ITEMS;
    static constexpr size_t TILE_BYTES = (size_t)TILE * sizeof(typename Tr::elem_t);
    static constexpr size_t RING_BYTES = 3 * TILE_BYTES;
    static constexpr size_t HIST_OFF = RING_BYTES;
    static constexpr size_t GPTR_OFF = HIST_OFF + (size_t)WARPS * RADIX * 4;  // [2][256] u64
    static constexpr size_t EXCL_OFF = GPTR_OFF + 2 * RADIX * 8;               // [2][256] u32
    static constexpr size_t CNT_OFF = EXCL_OFF + 2 * RADIX * 4;                // [2][256] u32
    static constexpr size_t MISC_OFF = CNT_OFF + 2 * RADIX * 4;                // [0..2] tickets, [4..11] scan totals
    static constexpr size_t MBAR_OFF = MISC_OFF + 64;                          // [3] mbarriers
    static constexpr size_t TOTAL = MBAR_OFF + 32;
};

// The hand-over barriers of the ring kernel come in two copies, used by even
// and odd tiles in turn: the look-back warps may finish tile i before the
// workers have waited for tile i-1 (and the workers may count tile i+1 before a
// slow look-back warp has picked up tile i), and a named barrier must never see
// the arrivals of two phases at once.
constexpr int RB_CNT = 2, RB_EXCL = 4, RB_GPTR = 6;  // + parity
template <int ID, int COUNT>
__device__ __forceinline__ void bar_sync_p(int parity) {
    if (parity) bar_sync<ID + 1, COUNT>();
    else bar_sync<ID, COUNT>();
}
template <int ID, int COUNT>
__device__ __forceinline__ void bar_arrive_p(int parity) {
    if (parity) bar_arrive<ID + 1, COUNT>();
    else bar_arrive<ID, COUNT>();
}

// One round of the look-back for one digit: LB_BATCH status words v[0] (nearest
// predecessor) .. v[LB_BATCH-1], evaluated without data-dependent branches.
// Returns 0 = all partial (sum them, go on), 1 = reached an inclusive prefix,
// 2 = met a word that is not published yet (ask again).  *part = the sum to add.
__device__ __forceinline__ int lookback_eval(const uint32_t (&v)[LB_BATCH], uint32_t* part) {
    uint32_t incl_mask = 0, unpub_mask = 0;
#pragma unroll
    for (int j = 0; j < LB_BATCH; ++j) {
        incl_mask |= (v[j] >> 31) << j;
        unpub_mask |= ((v[j] >> 30) == 0 ? 1u : 0u) << j;
    }
    const int first = incl_mask ? (__ffs(incl_mask) - 1) : (LB_BATCH - 1);  // last word that counts
    const uint32_t upto = (2u << first) - 1u;
    uint32_t sum = 0;
#pragma unroll
    for (int j = 0; j < LB_BATCH; ++j) sum += ((upto >> j) & 1u) ? (v[j] & LB_VALUE_MASK) : 0u;
    *part = sum;
    if (unpub_mask & upto) return 2;
    return incl_mask ? 1 : 0;
}

// look-back warps, one tile: LBT threads, NDL = 256 / LBT digits each, all digits
// of a thread advancing together, one batch per digit per L2 round trip
template <typename Tr, int WORKERS, int LBT>
__device__ __forceinline__ void ring_lookback(const PassArgs<Tr>& a, typename Tr::elem_t* dst, uint32_t tile, int parity,
                                              const uint32_t* cnt, const uint32_t* excl, unsigned long long* gptr,
                                              const unsigned long long (&bin)[RADIX / LBT]) {
    using E = typename Tr::elem_t;
    constexpr int NDL = RADIX / LBT;
    const int ltid = (int)threadIdx.x - WORKERS;
    uint32_t* lb = a.lookback + (size_t)tile * RADIX;
    bar_sync_p<RB_CNT, WORKERS + LBT>(parity);
    uint32_t sum[NDL];
    int pos[NDL];  // nearest predecessor not yet summed; < 0: closed
#pragma unroll
    for (int q = 0; q < NDL; ++q) { sum[q] = 0; pos[q] = (int)tile - 1; }
    const uint32_t* col = a.lookback + ltid;
    for (;;) {
        bool any = false;
#pragma unroll
        for (int q = 0; q < NDL; ++q) any |= pos[q] >= 0;
        if (!any) break;
        uint32_t v[NDL][LB_BATCH];
#pragma unroll
        for (int q = 0; q < NDL; ++q)
#pragma unroll
            for (int j = 0; j < LB_BATCH; ++j)
                v[q][j] = (pos[q] - j >= 0) ? ld_relaxed_u32(col + q * LBT + (size_t)(pos[q] - j) * RADIX) : LB_FLAG_INCLUSIVE;
#pragma unroll
        for (int q = 0; q < NDL; ++q) {
            if (pos[q] < 0) continue;
            uint32_t part;
            const int state = lookback_eval(v[q], &part);
            if (state == 2) continue;  // same batch again
            sum[q] += part;
            if (state == 1) {
                pos[q] = -1;
                st_relaxed_u32(lb + q * LBT + ltid, LB_FLAG_INCLUSIVE | (sum[q] + cnt[q * LBT + ltid]));
            } else {
                pos[q] -= LB_BATCH;
            }
        }
    }
    bar_sync_p<RB_EXCL, WORKERS + LBT>(parity);
#pragma unroll
    for (int q = 0; q < NDL; ++q) {
        const int d = q * LBT + ltid;
        const unsigned long long gbase = bin[q] + sum[q];
        gptr[d] = Tr::dst_base((unsigned)d, dst, a.prm) + (gbase - excl[d]) * sizeof(E);
        if (a.bins_out != nullptr && tile == a.num_tiles - 1) a.bins_out[d] = gbase + cnt[d];
    }
    bar_arrive_p<RB_GPTR, WORKERS + LBT>(parity);
}


template <typename Tr, int WORKERS, int ITEMS, int LBT, int MIN_BLOCKS>
__global__ void __launch_bounds__(WORKERS + LBT, MIN_BLOCKS)
onesweep_ring_kernel(const __grid_constant__ PassArgs<Tr> a) {
    using E = typename Tr::elem_t;
    using SM = RingSmem<Tr, WORKERS, ITEMS>;
    constexpr int WARPS = WORKERS / 32;
    constexpr int TILE = WORKERS * ITEMS;
    constexpr uint32_t TILE_BYTES = (uint32_t)SM::TILE_BYTES;
    static_assert(WORKERS >= RADIX && WORKERS % 32 == 0, "one worker thread per digit needed");
    static_assert(LBT >= 32 && LBT % 32 == 0 && RADIX % LBT == 0, "look-back warps own 256 / LBT digits per thread");
    static_assert(TILE_BYTES % 16 == 0, "TMA bulk copies move multiples of 16 bytes");

    const PassPlan pp = a.plan->pass[a.pass];
    if (pp.skip) return;

    extern __shared__ __align__(128) unsigned char smem_raw[];
    E* const ring = reinterpret_cast<E*>(smem_raw);
    uint32_t* const warp_hist = reinterpret_cast<uint32_t*>(smem_raw + SM::HIST_OFF);
    unsigned long long* const gptr2 = reinterpret_cast<unsigned long long*>(smem_raw + SM::GPTR_OFF);
    uint32_t* const excl2 = reinterpret_cast<uint32_t*>(smem_raw + SM::EXCL_OFF);
    uint32_t* const cnt2 = reinterpret_cast<uint32_t*>(smem_raw + SM::CNT_OFF);
    uint32_t* const misc = reinterpret_cast<uint32_t*>(smem_raw + SM::MISC_OFF);
    unsigned long long* const mbar = reinterpret_cast<unsigned long long*>(smem_raw + SM::MBAR_OFF);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t wbase = (uint32_t)warp * 32u * ITEMS + (uint32_t)lane;
    const int pass = a.pass;

    const E* __restrict__ src = (pp.src_is_tmp ? a.tmp : a.keys) + a.portion_off;
    E* __restrict__ dst = pp.src_is_tmp ? a.keys : a.tmp;
    const uint32_t full_tiles = a.portion_n / (uint32_t)TILE;  // tiles 0..full_tiles-1 are whole

    // thread 0 is the producer: draws tickets, starts the bulk copies (see the
    // two-slot kernel in sort_kernels.cuh for why a CTA holds at most two tickets)
    auto fetch = [&](int slot, uint32_t t) {
        misc[slot] = t;
        if (t < full_tiles) {
            mbar_expect_tx(&mbar[slot], TILE_BYTES);
            bulk_load(ring + (size_t)slot * TILE, src + (size_t)t * TILE, TILE_BYTES, &mbar[slot]);
        }
    };
    if (tid == 0) {
        mbar_init(&mbar[0], 1);
        mbar_init(&mbar[1], 1);
        mbar_init(&mbar[2], 1);
        fence_mbar_init();
        fetch(0, atomicAdd(a.ticket, 1u));
    }
    for (int i = tid; i < WARPS * RADIX; i += WORKERS + LBT) warp_hist[i] = 0;
    __syncthreads();

    if (tid >= WORKERS) {
        // ---- look-back warps
        constexpr int NDL = RADIX / LBT;
        unsigned long long bin[NDL];
#pragma unroll
        for (int q = 0; q < NDL; ++q) bin[q] = a.bins_in[q * LBT + (tid - WORKERS)];
        for (uint32_t it = 0;; ++it) {
            const uint32_t tile = misc[it % 3u];
            if (tile >= a.num_tiles) break;
            const int b = (int)(it & 1u);
            ring_lookback<Tr, WORKERS, LBT>(a, dst, tile, b, cnt2 + b * RADIX, excl2 + b * RADIX, gptr2 + b * RADIX, bin);
        }
        return;
    }

    // ---- workers.  View of the tile being written out (the previous one).
    PassShared<Tr, WORKERS, ITEMS> prev;
    prev.warp_hist = warp_hist;
    prev.misc = misc;
    prev.mbar = mbar;
    uint32_t prev_tile = 0, prev_n = 0;
    int prev_parity = 0;
    bool have_prev = false;

    // deferred write-out of the previous tile (needs the look-back warps' gptr)
    auto flush_prev = [&]() {
        bar_sync_p<RB_GPTR, WORKERS + LBT>(prev_parity);  // all workers past the scatter + gptr of the previous tile published
        bool bad = false;
        const bool verify = !(Tr::NEEDS_STABLE || pass == 0);
        if (prev_n == (uint32_t)TILE) {
            bad = verify ? write_out<Tr, WORKERS, ITEMS, true, true>(prev, a, pp.xf_out, prev_n, tid)
                         : write_out<Tr, WORKERS, ITEMS, true, false>(prev, a, pp.xf_out, prev_n, tid);
        } else {
            bad = verify ? write_out<Tr, WORKERS, ITEMS, false, true>(prev, a, pp.xf_out, prev_n, tid)
                         : write_out<Tr, WORKERS, ITEMS, false, false>(prev, a, pp.xf_out, prev_n, tid);
        }
        return bad;
    };

    uint32_t it = 0;
    for (;; ++it) {
        const int slot = (int)(it % 3u);
        const int b = (int)(it & 1u);
        const uint32_t tile = misc[slot];
        if (tile >= a.num_tiles) break;
        uint32_t next = 0;
        if (tid == 0) next = atomicAdd(a.ticket, 1u);
        E* const cur = ring + (size_t)slot * TILE;
        uint32_t* const cnt = cnt2 + b * RADIX;
        uint32_t* const excl = excl2 + b * RADIX;
        const E* src_tile = src + (size_t)tile * TILE;
        const bool full = tile < full_tiles;
        const uint32_t tile_n = full ? (uint32_t)TILE : a.portion_n - tile * (uint32_t)TILE;

        // ---- stage: keys of the tile into registers, warp-striped
        E e[ITEMS];
        if (full) {
            mbar_wait(&mbar[slot], (it / 3u) & 1u);
#pragma unroll
            for (int i = 0; i < ITEMS; ++i) e[i] = cur[wbase + i * 32];
        } else {
#pragma unroll
            for (int i = 0; i < ITEMS; ++i)
                if (wbase + i * 32 < tile_n) e[i] = ld_elem(src_tile + wbase + i * 32);
        }
        if (pp.xf_in) {
#pragma unroll
            for (int i = 0; i < ITEMS; ++i) e[i] = Tr::xform_in(e[i], a.prm);
        }

        // ---- 1. count
        uint32_t* my_hist = warp_hist + warp * RADIX;
        uint32_t rank[Tr::NEEDS_STABLE ? ITEMS : 1];
        if (Tr::NEEDS_STABLE) {
            const unsigned lane_lt = (1u << lane) - 1u;
#pragma unroll
            for (int i = 0; i < ITEMS; ++i)
                rank[Tr::NEEDS_STABLE ? i : 0] =
                    ballot_rank<false>(Tr::digit(e[i], pass, a.prm), wbase + i * 32 < tile_n, my_hist, lane, lane_lt);
        } else {
#pragma unroll
            for (int i = 0; i < ITEMS; ++i)
                if (full || wbase + i * 32 < tile_n) atomicAdd(&my_hist[hidx(Tr::digit(e[i], pass, a.prm))], 1u);
        }
        bar_sync<BAR_WORKERS, WORKERS>();
        // the slot two tiles back was written out during the last iteration: refill it
        if (tid == 0) fetch((int)((it + 1u) % 3u), next);

        // ---- 2. digit owners: tile count -> publish -> scan -> per-warp first slots
        uint32_t count = 0;
        uint32_t* lb = a.lookback + (size_t)tile * RADIX;
        if (tid < RADIX) {
            count = column_total<WARPS>(warp_hist, tid);
            st_relaxed_u32(lb + tid, (tile == 0 ? LB_FLAG_INCLUSIVE : LB_FLAG_PARTIAL) | count);
            cnt[tid] = count;
        }
        bar_arrive_p<RB_CNT, WORKERS + LBT>(b);
        uint32_t incl = count;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(FULL_MASK, incl, o);
            if (lane >= o) incl += y;
        }
        if (lane == 31 && warp < RADIX / 32) misc[4 + warp] = incl;
        bar_sync<BAR_WORKERS, WORKERS>();
        if (tid < RADIX) {
            uint32_t base = 0;
#pragma unroll
            for (int w = 0; w < RADIX / 32; ++w) base += (w < warp) ? misc[4 + w] : 0u;
            const uint32_t ex = base + incl - count;
            excl[tid] = ex;
            column_bases<WARPS>(warp_hist, tid, ex);
        }
        bar_arrive_p<RB_EXCL, WORKERS + LBT>(b);
        bar_sync<BAR_WORKERS, WORKERS>();

        // ---- 3. scatter into the slot the keys came from
#pragma unroll
        for (int i = 0; i < ITEMS; ++i)
            if (full || wbase + i * 32 < tile_n) {
                const unsigned d = Tr::digit(e[i], pass, a.prm);
                const uint32_t s = Tr::NEEDS_STABLE ? rank[Tr::NEEDS_STABLE ? i : 0] + my_hist[hidx(d)] : atomicAdd(&my_hist[hidx(d)], 1u);
                cur[s] = e[i];
            }

        // ---- 4. write out the PREVIOUS tile; its offsets were found while this one was ranked
        bool bad = false;
        if (have_prev) bad = flush_prev();
        else bar_sync<BAR_WORKERS, WORKERS>();  // first tile: only order the scatter before zeroing the counters
#pragma unroll
        for (int i = tid; i < WARPS * RADIX; i += WORKERS) warp_hist[i] = 0;
        fence_proxy_async();  // generic-proxy accesses of the slot just written out, before the next bulk copy into it
        if (bar_red_or<BAR_WORKERS, WORKERS>(bad)) {
            if (tid == 0) atomicAdd(a.stats, 1u);
            rerank_tile<Tr, WORKERS, ITEMS, LBT>(prev, &a, src + (size_t)prev_tile * TILE, prev_n, pp.xf_in, pp.xf_out);
            for (int i = tid; i < WARPS * RADIX; i += WORKERS) warp_hist[i] = 0;
            fence_proxy_async();
            bar_sync<BAR_WORKERS, WORKERS>();
        }
        prev.tile = cur;
        prev.gptr = gptr2 + b * RADIX;
        prev.excl = excl;
        prev.cnt = cnt;
        prev_tile = tile;
        prev_parity = b;
        prev_n = tile_n;
        have_prev = true;
    }
    // ---- drain: the last tile this CTA ranked
    if (have_prev) {
        const bool bad = flush_prev();
        if (bar_red_or<BAR_WORKERS, WORKERS>(bad)) {
            if (tid == 0) atomicAdd(a.stats, 1u);
            rerank_tile<Tr, WORKERS, ITEMS, LBT>(prev, &a, src + (size_t)prev_tile * TILE, prev_n, pp.xf_in, pp.xf_out);
        }
    }
}

}  // namespace b200
