// front_def.cuh -- definition of sort_front<FAM, W>: the launch sequence of one
// sort (replaces the body of RadixSort::radix_sort, reference mod.rs:61-176).
// Included only by inst_front.cu, which instantiates it for one (family, width).
// There is deliberately no CPU path: every call launches the sm_100a kernels or
// returns an error.
#pragma once
#include <cstdlib>
#include <cstring>

#include "sort_impl.cuh"

namespace b200 {

template <int FAM, int W, int I = 0>
cudaError_t launch_pass_idx(int variant, const PassArgs<typename FamilyTraits<FAM, W>::type>& a, cudaStream_t st) {
    if constexpr (I >= num_variants<FAM, W>()) {
        return cudaErrorInvalidValue;
    } else {
        if (variant == I) return VariantEntry<FAM, W, I>::launch(a, st);
        return launch_pass_idx<FAM, W, I + 1>(variant, a, st);
    }
}

template <int FAM, int W>
constexpr size_t narrow_bins() { return (FAM == FAM_KEYS && W <= 2) ? ((size_t)1 << (8 * W)) : 0; }

template <int FAM, int W>
size_t temp_bytes_front(size_t n) {
    using Tr = typename FamilyTraits<FAM, W>::type;
    // the maximum over all variants, so a sweep never outgrows the workspace
    size_t m = 0;
    for (int i = 0; i < num_variants<FAM, W>(); ++i) {
        const VariantDesc d = Variants<FAM, W>::v[i];
        const size_t t = make_layout(n ? n : 1, W, Tr::DIGITS, d.threads * d.items, FAM != FAM_PARTITION, narrow_bins<FAM, W>()).total;
        if (t > m) m = t;
    }
    return m;
}

template <int FAM, int W>
int make_params(const SortCall& c, typename FamilyTraits<FAM, W>::type::Params* prm) {
    using Tr = typename FamilyTraits<FAM, W>::type;
    typename Tr::Params p{};
    if constexpr (FAM == FAM_KEYS) {
        p.xf = c.xf;
    } else if constexpr (FAM == FAM_RECORDS) {
        p.key_shift = c.key_shift;
        p.xf = c.xf;
    } else {
        using U = typename Tr::elem_t;
        p.nsplit = c.nsplit;
        p.xf = c.xf;
        p.mask = c.bucket_mask;
        const U* raw = static_cast<const U*>(c.splitters);
        for (int j = 0; j < c.nsplit; ++j) p.spl[j] = to_ordered_rt<U>(raw[j], c.xf);
        for (int j = c.nsplit; j < MAX_SPLITTERS; ++j) std::memset(&p.spl[j], 0xFF, sizeof(U));  // padding: largest ordered value
        for (int j = 1; j < c.nsplit; ++j)
            if (ordered_less<U>(p.spl[j], p.spl[j - 1])) return fail(B200SORT_ERR_INVALID_ARG, "splitters are not ascending");
        if (c.peer_bases)
            for (int j = 0; j <= c.nsplit; ++j) {
                if (!c.peer_bases[j] || ((uintptr_t)c.peer_bases[j] % alignof(U)) != 0)
                    return fail(B200SORT_ERR_INVALID_ARG, "peer buffer null or not aligned to the element size");
                p.peer[j] = (unsigned long long)(uintptr_t)c.peer_bases[j];
            }
        if (c.phase == SortCall::PHASE_CLAIM) {
            if (!c.peer_bases || !c.claim_cursors || !c.claim_caps || !c.claim_flags)
                return fail(B200SORT_ERR_INVALID_ARG, "claim phase needs region bases, cursors, capacities and a flag word");
            p.claim = 1u;
            for (int j = 0; j <= c.nsplit; ++j) {
                if (!c.claim_cursors[j] || (c.claim_cursors[j] & 7)) return fail(B200SORT_ERR_INVALID_ARG, "cursor null or not 8-byte aligned");
                p.cursor[j] = c.claim_cursors[j];
                p.cap[j] = c.claim_caps[j];
            }
            p.flags = (unsigned long long)(uintptr_t)c.claim_flags;
        }
    }
    *prm = p;
    return B200SORT_OK;
}

template <int FAM, int W>
int sort_front(const SortCall& c) {
    using Tr = typename FamilyTraits<FAM, W>::type;
    using E = typename Tr::elem_t;
    static_assert(sizeof(E) == W, "element width");
    constexpr int D = Tr::DIGITS;
    const size_t n = c.n;
    cudaStream_t st = c.st;
    typename Tr::Params prm;
    if (int rc = make_params<FAM, W>(c, &prm)) return rc;
    const int ph = c.phase;
    // (a second scatter over the same counts zeroes the same region: the histogram in it is spent by then)
    const bool do_hist = ph == SortCall::PHASE_ALL || ph == SortCall::PHASE_COUNT;
    const bool do_zero = do_hist || (ph == SortCall::PHASE_SCATTER && c.rezero);
    const bool do_plan = do_hist;
    const bool claim = ph == SortCall::PHASE_CLAIM;
    const bool do_scatter = ph == SortCall::PHASE_ALL || ph == SortCall::PHASE_SCATTER || claim;
    const bool do_count = do_plan;  // the call that delivers the counts
    if (n == 0) {
        if (c.hist_copy && do_count) B200_CU(cudaMemsetAsync(c.hist_copy, 0, (size_t)c.hist_copy_words * 8, st));
        return B200SORT_OK;
    }
    E* d_elems = static_cast<E*>(c.elems);
    if (!d_elems || !c.ws) return fail(B200SORT_ERR_INVALID_ARG, "null device pointer");
    if (((uintptr_t)d_elems % alignof(E)) != 0 || (c.alt_out && ((uintptr_t)c.alt_out % alignof(E)) != 0))
        return fail(B200SORT_ERR_INVALID_ARG, "keys not aligned to the element size");
    if (((uintptr_t)c.ws % 256) != 0) return fail(B200SORT_ERR_INVALID_ARG, "workspace not 256-byte aligned");

    // small slices: the whole sort in one launch of one CTA (sort_kernels.cuh, small_sort_kernel)
    if constexpr (FAM != FAM_PARTITION) {
        if (ph == SortCall::PHASE_ALL && n <= (size_t)SmallSmem<Tr>::TILE && !c.out_ms && small_sort_enabled()) {
            auto k = small_sort_kernel<Tr>;
            constexpr size_t SS = SmallSmem<Tr>::TOTAL;
            static std::atomic<unsigned> small_configured{0};  // bit per device
            int dev = 0;
            cudaGetDevice(&dev);
            if (SS > 48 * 1024 && !(small_configured.load(std::memory_order_acquire) & (1u << (dev & 31)))) {
                B200_CU(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SS));
                small_configured.fetch_or(1u << (dev & 31), std::memory_order_release);
            }
            k<<<1, SMALL_THREADS, SS, st>>>(d_elems, (uint32_t)n, prm, (c.flags & B200SORT_FLAG_NO_SKIP) ? 1 : 0);
            count_launch();
            B200_CU(cudaGetLastError());
            return B200SORT_OK;
        }
    }

    int variant = variant_override(FAM, W);
    if (variant < 0 || variant >= num_variants<FAM, W>()) variant = 0;
    const VariantDesc vd = Variants<FAM, W>::v[variant];
    const int tile = vd.threads * vd.items;
    const Layout L = make_layout(n, W, D, tile, FAM != FAM_PARTITION, narrow_bins<FAM, W>());
    if (c.ws_bytes < L.total) return fail(B200SORT_ERR_INVALID_ARG, "workspace smaller than b200sort_temp_bytes()");

    unsigned char* ws = static_cast<unsigned char*>(c.ws);
    // key-only 1- and 2-byte keys: value histogram + fill (narrow_kernels.cuh; the reference's own
    // counting_sort, mod.rs:40-59, is the one-byte case).  FLAG_NO_SKIP and the timed entry point keep
    // the onesweep passes, so tests and sweeps still reach K3 for these widths.
    if constexpr (FAM == FAM_KEYS && W <= 2) {
        if (ph == SortCall::PHASE_ALL && !c.out_ms && !(c.flags & B200SORT_FLAG_NO_SKIP) && narrow_fill_enabled()) {
            using NC = NarrowCfg<E>;
            unsigned long long* vhist = reinterpret_cast<unsigned long long*>(ws + L.narrow_off);
            unsigned long long* voffs = vhist + NC::NV;
            int dev = 0, sms = 148;
            cudaGetDevice(&dev);
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            auto hk = hist_values_kernel<E>;
            static std::atomic<unsigned> configured{0};  // bit per device
            if (NC::SMEM > 48 * 1024 && !(configured.load(std::memory_order_acquire) & (1u << (dev & 31)))) {
                B200_CU(cudaFuncSetAttribute(hk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)NC::SMEM));
                configured.fetch_or(1u << (dev & 31), std::memory_order_release);
            }
            uint32_t* partial = reinterpret_cast<uint32_t*>(ws + L.narrow_partial_off);
            unsigned slabs;
            if constexpr (NC::SPLIT == 1) {  // u8: two CTAs per SM, one global atomic per bin and CTA
                B200_CU(cudaMemsetAsync(vhist, 0, (size_t)NC::NV * 8, st));
                const size_t vecs = (n * W + 15) / 16;
                const size_t want = (vecs + (size_t)NC::THREADS * 4 - 1) / ((size_t)NC::THREADS * 4);
                slabs = (unsigned)std::max<size_t>(1, std::min<size_t>(2 * (size_t)sms, want));
            } else {                         // u16: one CTA per SM, in pairs that share a slab of the input
                slabs = narrow_slabs16(n, sms);
            }
            hk<<<slabs * NC::SPLIT, NC::THREADS, NC::SMEM, st>>>(d_elems, n, vhist, partial, c.xf);
            count_launch();
            B200_CU(cudaGetLastError());
            if constexpr (NC::SPLIT == 2) {
                reduce_partials_kernel<<<NC::NV / 1024, 1024, 0, st>>>(partial, slabs, NC::NV, vhist);
                count_launch();
                B200_CU(cudaGetLastError());
            }
            scan_values_kernel<NC::NV><<<1, 1024, 0, st>>>(vhist, voffs);
            count_launch();
            B200_CU(cudaGetLastError());
            size_t fill_blocks = (n + 65535) / 65536;
            if (fill_blocks > (size_t)sms * 2) fill_blocks = (size_t)sms * 2;
            fill_values_kernel<E><<<(unsigned)fill_blocks, 1024, 0, st>>>(d_elems, n, voffs, c.xf);
            count_launch();
            B200_CU(cudaGetLastError());
            return B200SORT_OK;
        }
    }
    // alt_out: single-pass callers (the multi-GPU partition) name the destination
    // themselves; the result stays there and nothing is copied back
    if (FAM == FAM_PARTITION && !c.alt_out) return fail(B200SORT_ERR_INVALID_ARG, "partition needs a destination");
    E* tmp = c.alt_out ? static_cast<E*>(c.alt_out) : reinterpret_cast<E*>(ws + L.tmp_off);
    unsigned long long* bins = reinterpret_cast<unsigned long long*>(ws + L.bins_off);
    Plan* plan = reinterpret_cast<Plan*>(ws + L.plan_off);
    unsigned long long* hist = reinterpret_cast<unsigned long long*>(ws + L.hist_off);
    uint32_t* tickets = reinterpret_cast<uint32_t*>(ws + L.ticket_off);
    uint32_t* lookback = reinterpret_cast<uint32_t*>(ws + L.lb_off);
    uint32_t* stats = reinterpret_cast<uint32_t*>(ws + L.stats_off);

    if (ph == SortCall::PHASE_READ_STATS) {
        B200_CU(cudaStreamSynchronize(st));
        B200_CU(cudaMemcpy(c.out_stats, stats, sizeof(unsigned), cudaMemcpyDeviceToHost));
        return B200SORT_OK;
    }
    PhaseTimer timer(c.out_ms != nullptr, st);
    timer.mark();
    if (do_zero) B200_CU(cudaMemsetAsync(ws + L.zero_off, 0, L.zero_bytes, st));
    if (claim) {  // no histogram, no plan kernel, no look-back: an all-zero plan (pass 0 executes, reads the input) and fresh tickets
        B200_CU(cudaMemsetAsync(ws + L.plan_off, 0, sizeof(Plan), st));
        B200_CU(cudaMemsetAsync(ws + L.ticket_off, 0, L.lb_off - L.ticket_off, st));
    }

    // K1: one read, all digit histograms
    if (do_hist) {
        const E* h_elems = d_elems;
        const size_t h_n = n;
        int dev = 0, sms = 148;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        constexpr int HT = HistCfg<Tr>::THREADS;
        constexpr size_t HS = HistCfg<Tr>::SMEM;
        auto hk = hist_all_digits_kernel<Tr, HT>;
        static std::atomic<unsigned> configured{0};  // bit per device
        if (HS > 48 * 1024 && !(configured.load(std::memory_order_acquire) & (1u << (dev & 31)))) {
            B200_CU(cudaFuncSetAttribute(hk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)HS));
            configured.fetch_or(1u << (dev & 31), std::memory_order_release);
        }
        const size_t per_sm = std::max<size_t>(1, std::min<size_t>(2048 / HT, (220 * 1024) / HS));
        const size_t vecs = (h_n * W + 15) / 16;
        size_t blocks = (vecs + (size_t)HT * 4 - 1) / ((size_t)HT * 4);
        const size_t cap = (size_t)sms * per_sm;
        if (blocks > cap) blocks = cap;
        if (blocks < 1) blocks = 1;
        if (h_n) {
            hk<<<(unsigned)blocks, HT, HS, st>>>(h_elems, h_n, hist, prm);
            count_launch();
            B200_CU(cudaGetLastError());
        }
    }
    timer.mark();
    if (do_plan) {
        if (c.hist_copy)
            B200_CU(cudaMemcpyAsync(c.hist_copy, hist, (size_t)c.hist_copy_words * 8, cudaMemcpyDeviceToDevice, st));
        // K2: bucket bases + pass plan
        scan_plan_kernel<<<1, RADIX, 0, st>>>(hist, bins, plan, (unsigned long long)n, D, L.portions,
                                              (c.flags & B200SORT_FLAG_NO_SKIP) ? 1 : 0);
        count_launch();
        B200_CU(cudaGetLastError());
    }
    timer.mark();
    if (!do_scatter) return B200SORT_OK;
    // fused exchange: bucket r starts at element peer_offsets[r] of rank r's receive buffer
    if (c.peer_offsets && !claim)
        B200_CU(cudaMemcpyAsync(bins, c.peer_offsets, (size_t)(c.nsplit + 1) * 8, cudaMemcpyDeviceToDevice, st));

    // K3: one launch per digit per portion
    for (int p = 0; p < D; ++p) {
        size_t lb_cursor = 0;  // tiles of the portions before q: every pass lays its portions out alike
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
            a.grid = a.num_tiles;
            a.bins_in = bins + ((size_t)p * L.portions + q) * RADIX;
            a.bins_out = (q + 1 < L.portions) ? bins + ((size_t)p * L.portions + q + 1) * RADIX : nullptr;
            a.lookback = lookback + lb_cursor * RADIX;
            a.lookback_half = (FAM == FAM_PARTITION) ? 0 : L.tiles_total * RADIX;
            a.ticket = tickets + (size_t)p * L.portions + q;
            a.stats = stats;
            a.force_rerank = force_rerank() ? 1u : 0u;
            a.opaque_zero = 0u;
            a.prm = prm;
            lb_cursor += a.num_tiles;
            B200_CU((launch_pass_idx<FAM, W>(variant, a, st)));
        }
        timer.mark();
    }
    // result in tmp after an odd number of executed passes (mod.rs:170-174)
    if (!c.alt_out) {
        const size_t bytes = n * W;
        size_t blocks = (bytes / 16 + 255) / 256;
        if (blocks > 148 * 8) blocks = 148 * 8;
        if (blocks < 1) blocks = 1;
        copy_back_kernel<E><<<(unsigned)blocks, 256, 0, st>>>(d_elems, tmp, n, plan);
        count_launch();
        B200_CU(cudaGetLastError());
    }
    timer.mark();

    if (c.out_ms) {
        if (c.out_len < D + 4) return fail(B200SORT_ERR_INVALID_ARG, "out_len < digits + 4");
        B200_CU(cudaStreamSynchronize(st));
        // marks: 0 start | 1 after memset+K1 | 2 after K2 | 3..2+D after each pass | 3+D after copy-back
        for (int i = 0; i < D + 3; ++i) cudaEventElapsedTime(&c.out_ms[i], timer.ev[i], timer.ev[i + 1]);
        cudaEventElapsedTime(&c.out_ms[D + 3], timer.ev.front(), timer.ev.back());
    }
    return B200SORT_OK;
}

}  // namespace b200
