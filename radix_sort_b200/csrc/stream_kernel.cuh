// stream_kernel.cuh -- K3, streaming form (kernel id 3): resident CTAs run the
// one-tile code in a loop and keep the NEXT tile's keys in flight while the
// current tile waits on its look-back and is written out.
//
// Why (r01 sweep v10, timing ablations of the one-tile kernel at 2^28 u64 keys):
// load 0.46 ms + rank/scatter 0.22 + look-back 0.12 + write-out 0.40 add up to
// the whole 1.2 ms pass -- with two CTAs per SM the phases of a tile form one
// serial chain and hardly overlap, and HBM sees loads only in bursts.  The key
// registers e[] are dead once the tile is scattered into shared memory, so the
// loads of the next tile can be issued into the SAME registers right there, at
// no register or shared-memory cost, and fly during the look-back, the barrier
// and the write-out.  The ticket of the next tile is drawn one phase earlier
// still (top of the loop), so its round trip to L2 is hidden too.
//
// Forward progress: a CTA holds at most the tile it works on and one prefetched
// ticket.  The smallest unfinished tile id is either being worked on (all its
// predecessors are finished, so its look-back ends) or is the prefetched ticket
// of a CTA whose current tile has a smaller id -- which is finished by
// assumption, so that CTA moves on.  No tile waits on a tile that cannot start.
#pragma once
#include "sort_kernels.cuh"

namespace b200 {

template <typename Tr, int THREADS, int ITEMS, int MIN_BLOCKS, int LBB, int ABL = 0>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS)
onesweep_stream_kernel(const __grid_constant__ PassArgs<Tr> a) {
    using E = typename Tr::elem_t;
    using SM = PassSmem<Tr, THREADS, ITEMS>;
    constexpr int WARPS = THREADS / 32;
    constexpr int TILE = THREADS * ITEMS;
    static_assert(THREADS >= RADIX && THREADS % 32 == 0, "one thread per digit needed");

    uint32_t ticket = 0;
    if (threadIdx.x == 0) ticket = atomicAdd(a.ticket, 1u);
    const PassPlan pp = a.plan->pass[a.pass];
    if (pp.skip) return;

    extern __shared__ __align__(128) unsigned char smem_raw[];
    PassShared<Tr, THREADS, ITEMS> sh(smem_raw, SM::TILE_BYTES);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t wbase = (uint32_t)warp * 32u * ITEMS + (uint32_t)lane;
    uint32_t* my_hist = sh.warp_hist + warp * RADIX;

    const E* __restrict__ src = (pp.src_is_tmp ? a.tmp : a.keys) + a.portion_off;
    E* __restrict__ dst = pp.src_is_tmp ? a.keys : a.tmp;

    if (tid == 0) sh.misc[0] = ticket;
#pragma unroll
    for (int i = lane; i < RADIX; i += 32) my_hist[i] = 0;
    __syncthreads();
    uint32_t tile = sh.misc[0];
    if (tile >= a.num_tiles) return;

    // warp-striped loads of one tile into e[]; slots past the end stay untouched
    E e[ITEMS];
    auto load_tile = [&](uint32_t t) {
        const E* p = src + (size_t)t * TILE + wbase;
        const uint32_t n = a.portion_n - t * (uint32_t)TILE;  // >= TILE for every tile but the last
        if (n >= (uint32_t)TILE) {
#pragma unroll
            for (int i = 0; i < ITEMS; ++i) e[i] = ld_elem(p + i * 32);
        } else {
#pragma unroll
            for (int i = 0; i < ITEMS; ++i)
                if (wbase + i * 32 < n) e[i] = ld_elem(p + i * 32);
        }
    };
    load_tile(tile);

    for (uint32_t it = 0;; ++it) {
        // ticket of the tile after this one: asked for now, parked in shared memory
        // after the count phase, read by everybody after the scatter
        uint32_t drawn = 0;
        if (tid == 0) drawn = atomicAdd(a.ticket, 1u);
        const int nslot = (int)((it + 1u) & 1u);
        auto park_ticket = [&]() {
            if (tid == 0) sh.misc[nslot] = drawn;
        };
        auto prefetch = [&]() {
            // my warp's counters are mine alone from here to the next tile's count
#pragma unroll
            for (int i = lane; i < RADIX; i += 32) my_hist[i] = 0;
            const uint32_t nt = sh.misc[nslot];
            if (nt < a.num_tiles) load_tile(nt);
        };
        const E* src_tile = src + (size_t)tile * TILE;
        const uint32_t tile_n = min((uint32_t)TILE, a.portion_n - tile * (uint32_t)TILE);
        bool bad;
        if (tile_n == (uint32_t)TILE)
            bad = rank_and_scatter<Tr, THREADS, ITEMS, true, LBB, ABL>(sh, a, pp, e, dst, tile, tile_n, park_ticket, prefetch);
        else
            bad = rank_and_scatter<Tr, THREADS, ITEMS, false, LBB, ABL>(sh, a, pp, e, dst, tile, tile_n, park_ticket, prefetch);
        // end-of-tile barrier: the write-out has read sh.tile / sh.gptr before the
        // next tile's phases overwrite them
        if (!Tr::NEEDS_STABLE && a.pass == 0 && !a.force_rerank) {
            __syncthreads();
        } else if (__syncthreads_or(bad || a.force_rerank)) {
            if (tid == 0) atomicAdd(a.stats, 1u);
            rerank_tile<Tr, THREADS, ITEMS>(sh, &a, src_tile, tile_n, pp.xf_in, pp.xf_out);
#pragma unroll
            for (int i = lane; i < RADIX; i += 32) my_hist[i] = 0;
            __syncthreads();
        }
        tile = sh.misc[nslot];
        if (tile >= a.num_tiles) return;
    }
}

}  // namespace b200
