// narrow_kernels.cuh -- key-only sorts of 1- and 2-byte keys without any ranking.
//
// Equal keys are bit-identical (radix_digits.rs:7-21,55-69: the ordered image is a
// bijection of the bit pattern), so the sorted slice is a function of the value
// histogram alone: value v, count[v] times, in get_digit's order.  That is the
// reference's own counting_sort (mod.rs:40-59: histogram, prefix, fill) taken to 16
// bits: one read of the keys (K1n) + one streaming write (K3n) = 2 s bytes per key
// instead of s (1 + 2 D) through the onesweep passes, and no shared-memory ranking.
//
//   K1n  hist_values_kernel   full histogram over all 2^(8s) values of the ordered key
//          u8 : 256 bins, every bin replicated 32 times (lane-private banks), like K1
//          u16: 65536 u32 bins are 256 KB -- more than one SM has.  Two CTAs share a
//               slab of the input, each counts the values whose top bit is its own
//               (32768 bins = 128 KB of shared memory each); the slab is read twice,
//               the second time mostly from L2.  Per-slab counts are stored, then summed
//               by reduce_partials_kernel.
//   K2n  scan_values_kernel   exclusive prefix over the bins -> offs[NV + 1]
//   K3n  fill_values_kernel   CTA b owns output elements [b C, (b + 1) C): finds its first
//          value by a two-level search of offs, then walks the values: short runs are
//          written by one thread each, middle ones by a warp, long ones by the whole CTA,
//          with 16-byte stores.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "key_traits.cuh"

namespace b200 {

template <typename U> struct NarrowCfg;
template <> struct NarrowCfg<uint8_t> {
    static constexpr int NV = 256, THREADS = 1024, R = 32, SPLIT = 1;
    static constexpr size_t SMEM = (size_t)NV * R * 4;
};
template <> struct NarrowCfg<uint16_t> {
    static constexpr int NV = 65536, THREADS = 1024, R = 1, SPLIT = 2;
    static constexpr size_t SMEM = (size_t)(NV / SPLIT) * 4;
};

constexpr int NARROW_MAX_SLABS = 80;  // u16: slabs of the input (two CTAs each); B200: 148 SMs -> 74

// slabs the u16 histogram is taken over for n keys on a device with `sms` SMs (host and layout agree on it)
inline unsigned narrow_slabs16(size_t n, int sms) {
    const size_t vecs = (n * 2 + 15) / 16;
    size_t want = (vecs + 4095) / 4096;
    if (want < 1) want = 1;
    size_t cap = (size_t)(sms / 2 < 1 ? 1 : sms / 2);
    if (cap > (size_t)NARROW_MAX_SLABS) cap = NARROW_MAX_SLABS;
    return (unsigned)(want < cap ? want : cap);
}

// K1n.  grid = SPLIT * slabs; CTA (slab, half) counts the keys of its slab whose ordered
// value falls into its half of the value range.
//   u8 : hist = NV u64 bins, zeroed, bumped with one global atomic per bin and CTA.
//   u16: `partial` = [slab][NV] u32, every CTA stores its 32768 counts with plain coalesced stores
//        (148 CTAs x 32768 u64 atomics on the same 65536 addresses took 0.2 ms -- more than the
//        whole histogram of 2^27 keys); reduce_partials_kernel adds the slabs up.
template <typename U>
__global__ void __launch_bounds__(NarrowCfg<U>::THREADS)
hist_values_kernel(const U* __restrict__ in, size_t n, unsigned long long* __restrict__ hist,
                   uint32_t* __restrict__ partial, int xf) {
    using C = NarrowCfg<U>;
    constexpr int VEC = 16 / (int)sizeof(U);
    extern __shared__ __align__(16) uint32_t bins[];
    const int tid = threadIdx.x;
    constexpr int NB = C::NV / C::SPLIT * C::R;
    for (int i = tid; i < NB; i += C::THREADS) bins[i] = 0;
    __syncthreads();
    const unsigned half = blockIdx.x % C::SPLIT;
    const unsigned slab = blockIdx.x / C::SPLIT, slabs = gridDim.x / C::SPLIT;
    auto count = [&](U raw) {
        const unsigned v = to_ordered_rt<U>(raw, xf);
        if constexpr (C::SPLIT == 1) {
            atomicAdd(&bins[v * C::R + (tid & (C::R - 1))], 1u);
        } else {
            if ((v >> 15) == half) atomicAdd(&bins[v & 0x7FFFu], 1u);
        }
    };
    // The vector loop works on whole 32-bit words: the ordered image of a 1- or 2-byte key is the key with its
    // sign bit flipped or not (no float kinds at these widths), i.e. ONE xor per word with a mask picked before
    // the loop, instead of a select per key (r03 ncu digest of the per-key form: issue-bound, 12 ALU instructions
    // per ATOMS).  u16 folds "is this value in my half" into the same xor: afterwards a key is mine iff its top bit is set.
    uint32_t wmask;
    if constexpr (sizeof(U) == 1) wmask = (xf != XF_NONE) ? 0x80808080u : 0u;
    else wmask = ((xf != XF_NONE) ? 0x80008000u : 0u) ^ (half ? 0u : 0x80008000u);
    // byte offsets, not indices: (value << 7 | lane << 2) comes out of one shift and one logic op per key, and the
    // array base rides in the atomic's uniform-register operand
    unsigned char* const bins_b = reinterpret_cast<unsigned char*>(bins);
    const uint32_t lane_off = (C::SPLIT == 1) ? (uint32_t)(tid & (C::R - 1)) * 4u : 0u;
    auto count_word = [&](uint32_t w) {
        w ^= wmask;
        if constexpr (sizeof(U) == 1) {
            static_assert(sizeof(U) != 1 || C::R == 32, "u8: value * 128 + lane * 4");
            atomicAdd(reinterpret_cast<uint32_t*>(bins_b + (((w << 7) & 0x7F80u) | lane_off)), 1u);
            atomicAdd(reinterpret_cast<uint32_t*>(bins_b + (((w >> 1) & 0x7F80u) | lane_off)), 1u);
            atomicAdd(reinterpret_cast<uint32_t*>(bins_b + (((w >> 9) & 0x7F80u) | lane_off)), 1u);
            atomicAdd(reinterpret_cast<uint32_t*>(bins_b + (((w >> 17) & 0x7F80u) | lane_off)), 1u);
        } else {
            // (a predicated `red.shared` does not help: ptxas turns it back into the same branch + warp-aggregated
            // ATOMS.POPC.INC, and an unconditional add of 0 / 1 would double the shared-memory wavefronts)
            if (w & 0x8000u) atomicAdd(reinterpret_cast<uint32_t*>(bins_b + ((w << 2) & 0x1FFFCu)), 1u);
            if (w & 0x80000000u) atomicAdd(reinterpret_cast<uint32_t*>(bins_b + ((w >> 14) & 0x1FFFCu)), 1u);
        }
    };
    auto count_vec = [&](uint4 q) {
        count_word(q.x); count_word(q.y); count_word(q.z); count_word(q.w);
    };
    const uintptr_t addr = (uintptr_t)in;
    size_t head = ((16 - (addr & 15)) & 15) / sizeof(U);
    if (head > n) head = n;
    const uint4* vec = reinterpret_cast<const uint4*>(in + head);
    const size_t nvec = (n - head) / VEC;
    const size_t tail = head + nvec * VEC;
    // slab = a contiguous run of vectors, so that the two CTAs of a slab walk the same lines
    const size_t per = (nvec + slabs - 1) / slabs;
    const size_t lo = (size_t)slab * per, hi = lo + per < nvec ? lo + per : nvec;
    size_t i = lo + tid;
    for (; i + 3 * (size_t)C::THREADS < hi; i += 4 * (size_t)C::THREADS) {
        const uint4 a = __ldg(vec + i), b = __ldg(vec + i + C::THREADS), c = __ldg(vec + i + 2 * C::THREADS),
                    d = __ldg(vec + i + 3 * C::THREADS);
        count_vec(a); count_vec(b); count_vec(c); count_vec(d);
    }
    for (; i < hi; i += C::THREADS) count_vec(__ldg(vec + i));
    if (slab == 0) {
        for (size_t j = tid; j < head; j += C::THREADS) count(in[j]);
        for (size_t j = tail + tid; j < n; j += C::THREADS) count(in[j]);
    }
    __syncthreads();
    if constexpr (C::SPLIT == 1) {
        for (int k = tid; k < C::NV; k += C::THREADS) {
            uint32_t c = 0;
#pragma unroll
            for (int r = 0; r < C::R; ++r) c += bins[k * C::R + ((r + tid) & (C::R - 1))];
            if (c) atomicAdd(&hist[k], (unsigned long long)c);
        }
    } else {
        constexpr int HB = C::NV / C::SPLIT;
        uint4* dst = reinterpret_cast<uint4*>(partial + (size_t)slab * C::NV + (size_t)half * HB);
        const uint4* src = reinterpret_cast<const uint4*>(bins);
        for (int j = tid; j < HB / 4; j += C::THREADS) dst[j] = src[j];
    }
}

// u16: hist[v] = sum over the slabs of partial[slab][v]; one thread per value, loads coalesced.
static __global__ void __launch_bounds__(1024)
reduce_partials_kernel(const uint32_t* __restrict__ partial, unsigned slabs, int nv, unsigned long long* __restrict__ hist) {
    const int v = blockIdx.x * 1024 + threadIdx.x;
    if (v >= nv) return;
    unsigned long long c = 0;
    for (unsigned s = 0; s < slabs; ++s) c += partial[(size_t)s * nv + v];
    hist[v] = c;
}

// K2n: one CTA of 1024 threads; offs[v] = keys with an ordered value below v, offs[NV] = n.
// Warp w owns the NV / 32 consecutive bins [w NV/32, (w+1) NV/32) and walks them 32 at a time
// (coalesced loads, shuffle scan, running carry); the warps' totals meet in shared memory.
template <int NV>
__global__ void __launch_bounds__(1024)
scan_values_kernel(const unsigned long long* __restrict__ hist, unsigned long long* __restrict__ offs) {
    constexpr int PER_WARP = NV / 32;              // 8 (u8) or 2048 (u16)
    constexpr int ROUNDS = PER_WARP >= 32 ? PER_WARP / 32 : 1;
    constexpr int LANES = PER_WARP >= 32 ? 32 : PER_WARP;
    __shared__ unsigned long long warp_tot[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned long long* h = hist + (size_t)warp * PER_WARP;
    unsigned long long tot = 0;
    for (int r = 0; r < ROUNDS; ++r) tot += lane < LANES ? h[r * 32 + lane] : 0ull;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xFFFFFFFFu, tot, o);
    if (lane == 0) warp_tot[warp] = tot;
    __syncthreads();
    unsigned long long carry = 0;
    for (int w = 0; w < warp; ++w) carry += warp_tot[w];
    unsigned long long* o_ = offs + (size_t)warp * PER_WARP;
    for (int r = 0; r < ROUNDS; ++r) {
        const unsigned long long c = lane < LANES ? h[r * 32 + lane] : 0ull;
        unsigned long long x = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long y = __shfl_up_sync(0xFFFFFFFFu, x, o);
            if (lane >= o) x += y;
        }
        if (lane < LANES) o_[r * 32 + lane] = carry + x - c;
        carry += __shfl_sync(0xFFFFFFFFu, x, 31);
    }
    if (tid == 1023) offs[NV] = carry;
}

// K3n.  out may be the buffer the keys were read from (the histogram has consumed them).
// Runs of one value inside the CTA's chunk come in three sizes: short ones are written by the thread
// that found them, middle ones by one warp each (32 lanes x 16 bytes per step; a u16 sort of 2^27
// uniform keys is 65536 runs of 4 KB: one CTA-wide loop per run left 3/4 of the CTA idle and was 0.2 ms
// of pure latency), long ones by the whole CTA.
template <typename U>
__device__ __forceinline__ void fill_run(U* __restrict__ out, unsigned long long s2, unsigned long long e2, U val, unsigned t, unsigned nt) {
    constexpr int VEC = 16 / (int)sizeof(U);
    const uintptr_t a0 = (uintptr_t)(out + s2);
    unsigned long long hd = ((16 - (a0 & 15)) & 15) / sizeof(U);  // head up to 16-byte alignment, vectors, tail
    if (hd > e2 - s2) hd = e2 - s2;
    const unsigned long long nv = (e2 - s2 - hd) / VEC;
    const unsigned long long tl = s2 + hd + nv * VEC;
    for (unsigned long long i = t; i < hd; i += nt) out[s2 + i] = val;
    const uint32_t w = sizeof(U) == 1 ? 0x01010101u * (uint32_t)val : 0x00010001u * (uint32_t)val;
    const uint4 vv = make_uint4(w, w, w, w);
    uint4* vp = reinterpret_cast<uint4*>(out + s2 + hd);
    for (unsigned long long i = t; i < nv; i += nt) vp[i] = vv;
    for (unsigned long long i = tl + t; i < e2; i += nt) out[i] = val;
}

template <typename U>
__global__ void __launch_bounds__(1024)
fill_values_kernel(U* __restrict__ out, size_t n, const unsigned long long* __restrict__ offs, int xf) {
    using C = NarrowCfg<U>;
    constexpr int NV = C::NV;
    constexpr int THREADS = 1024;
    constexpr unsigned SHORT_RUN = 32;      // elements below which the finding thread writes the run itself
    constexpr unsigned LONG_RUN = 32768;    // elements from which the whole CTA writes the run
    constexpr int QCAP = 1024;
    __shared__ unsigned s_first;
    __shared__ unsigned s_qn[2];            // [0] warp-sized runs, [1] CTA-sized runs
    __shared__ unsigned q_val[2][QCAP];
    __shared__ unsigned long long q_lo[2][QCAP], q_hi[2][QCAP];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // chunk of the output: whole 4096-element granules
    const size_t chunk = ((n + gridDim.x - 1) / gridDim.x + 4095) / 4096 * 4096;
    const size_t lo = (size_t)blockIdx.x * chunk;
    if (lo >= n) return;
    const size_t hi = lo + chunk < n ? lo + chunk : n;
    // first value of the chunk: the largest v with offs[v] <= lo.  Level 1: 1024 probes NV / 1024 apart,
    // level 2: the bracket the last passing probe opens.
    if (tid == 0) s_first = 0;
    __syncthreads();
    constexpr int STRIDE = NV >= THREADS ? NV / THREADS : 1;
    if (tid * STRIDE < NV && offs[tid * STRIDE] <= lo) atomicMax(&s_first, (unsigned)(tid * STRIDE));
    __syncthreads();
    if (STRIDE > 1) {
        const unsigned b0 = s_first;
        __syncthreads();
        if (tid < STRIDE && offs[b0 + tid] <= lo) atomicMax(&s_first, b0 + (unsigned)tid);
        __syncthreads();
    }
    unsigned vcur = s_first;
    for (;;) {
        if (tid < 2) s_qn[tid] = 0;
        __syncthreads();
        const unsigned v = vcur + (unsigned)tid;
        unsigned long long s = 0, e = 0;
        if (v < (unsigned)NV) {
            s = offs[v];
            e = offs[v + 1];
            if (s < lo) s = lo;
            if (e > hi) e = hi;
        }
        if (v < (unsigned)NV && e > s) {
            const U val = from_ordered_rt<U>((U)v, xf);
            if (e - s < SHORT_RUN) {
                for (unsigned long long p = s; p < e; ++p) out[p] = val;
            } else {
                const int q = e - s >= LONG_RUN ? 1 : 0;
                const unsigned slot = atomicAdd(&s_qn[q], 1u);
                q_val[q][slot] = val;
                q_lo[q][slot] = s;
                q_hi[q][slot] = e;
            }
        }
        // done when the last value of this batch already ends at or past the chunk's end
        const unsigned vlast = vcur + THREADS - 1 < (unsigned)NV - 1 ? vcur + THREADS - 1 : (unsigned)NV - 1;
        const bool done = offs[vlast + 1] >= hi || vcur + THREADS >= (unsigned)NV;
        __syncthreads();
        const unsigned qw = s_qn[0], qc = s_qn[1];
        for (unsigned k = warp; k < qw; k += THREADS / 32) fill_run<U>(out, q_lo[0][k], q_hi[0][k], (U)q_val[0][k], lane, 32);
        for (unsigned k = 0; k < qc; ++k) fill_run<U>(out, q_lo[1][k], q_hi[1][k], (U)q_val[1][k], tid, THREADS);
        if (done) return;
        vcur += THREADS;
        __syncthreads();
    }
}

}  // namespace b200
