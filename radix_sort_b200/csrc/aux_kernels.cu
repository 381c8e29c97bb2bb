// aux_kernels.cu -- K5 synthetic-input generators and K6 on-device verification.
//
// K5 restates the distribution families of the reference's src/distr.rs as
// counter-based generators: key i of a stream is a pure function of (seed, i),
// u = mix64(seed ^ mix64(i)), so a shard of a multi-GPU input is generated in
// place from its global indices and the multiset does not depend on the GPU
// count.  The reference seeds nothing (main.rs:27, tests.rs:28), so stream
// equality with the `rand` crate is neither possible nor needed.
//
// K6: sortedness under get_digit's order plus an order-independent multiset
// fingerprint.  For key-only sorts "sorted + same multiset" <=> bit-identical to
// the reference's output, which is how full-size runs are checked
// without hauling 8-16 GB through the host.
#include <cuda_runtime.h>
#include <math.h>

#include <string>

#include "../../include/b200sort.h"
#include "key_traits.cuh"

using namespace b200;

namespace b200 { void set_last_error(const std::string& msg); }  // b200sort_api.cu

namespace {

struct GenParams {
    int distr;
    double param;
    unsigned long long seed;
    unsigned long long index_base;
};

__device__ __forceinline__ unsigned long long draw(unsigned long long seed, unsigned long long i, unsigned long long j) {
    return mix64((seed ^ mix64(i)) + j * 0xD1B54A32D192ED03ull);
}

// Geometric(p): number of failures before the first success, k = floor(ln U / ln(1-p)),
// U in (0,1].  Family of MyExp (distr.rs:4-20).
__device__ __forceinline__ unsigned long long gen_geometric(unsigned long long u, double p) {
    const double U = ((double)(u >> 11) + 1.0) * (1.0 / 9007199254740992.0);
    const double k = floor(log(U) / log1p(-p));
    return k >= 18446744073709551615.0 ? ~0ull : (unsigned long long)k;
}

// Zipf(n, s) by rejection-inversion, the published algorithm rand_distr 0.4.3
// (Cargo.lock:144-151) implements; f32 arithmetic as ZipfU32/ZipfU64 use
// Zipf<f32> (distr.rs:55,109).  Result in [1, n].
__device__ __forceinline__ float zipf_inv_cdf(float p, float s, float q, float t) {
    const float pt = p * t;
    if (pt <= 1.0f) return pt;
    if (s != 1.0f) return powf(pt * q + s, 1.0f / q);
    return expf(pt - 1.0f);
}
__device__ __forceinline__ float gen_zipf(unsigned long long seed, unsigned long long i, float n, float s) {
    const float q = (s != 1.0f) ? 1.0f - s : 0.0f;
    const float t = (s != 1.0f) ? (powf(n, q) - s) / q : 1.0f + logf(n);
    for (unsigned long long j = 0; j < 64; ++j) {
        const unsigned long long u = draw(seed, i, j);
        const float p = (float)(u >> 40) * (1.0f / 16777216.0f);                 // [0,1)
        const float y = (float)((u >> 16) & 0xFFFFFFull) * (1.0f / 16777216.0f);  // [0,1)
        const float inv_b = zipf_inv_cdf(p, s, q, t);
        const float x = floorf(inv_b + 1.0f);
        float ratio = powf(x, -s);
        if (x > 1.0f) ratio *= powf(inv_b, s);
        if (y < ratio) return x;
    }
    return 1.0f;
}

template <typename U>
__device__ __forceinline__ U make_key(unsigned long long lo, unsigned long long hi) {
    (void)hi;
    return (U)lo;
}
template <>
__device__ __forceinline__ u128 make_key<u128>(unsigned long long lo, unsigned long long hi) {
    u128 k;
    k.lo = lo;
    k.hi = hi;
    return k;
}

// max value of the unsigned type as wide as the key, saturated to 64 bits
template <typename U>
__device__ __forceinline__ unsigned long long key_max64() {
    return sizeof(U) >= 8 ? ~0ull : ((1ull << (8 * (sizeof(U) & 7))) - 1ull);
}

template <typename U, bool IS_FLOAT>
__device__ __forceinline__ U gen_one(const GenParams g, unsigned long long i) {
    const unsigned long long u = draw(g.seed, i, 0);
    constexpr int BITS = (int)sizeof(U) * 8;
    const unsigned long long vmax = key_max64<U>();
    switch (g.distr) {
    case B200SORT_DISTR_GEOMETRIC: {
        unsigned long long k = gen_geometric(u, g.param);
        return make_key<U>(k, 0);  // truncating `as`, distr.rs:18
    }
    case B200SORT_DISTR_ZIPF: {
        // Zipf::new(u32::MAX as u64, s) / (u64::MAX, s); `as u32` saturates (distr.rs:62,68)
        const float n = sizeof(U) >= 8 ? 18446744073709551615.0f : (float)vmax;
        const float x = gen_zipf(g.seed, i, n, (float)g.param);
        unsigned long long k = x >= 18446744073709551615.0f ? ~0ull : (unsigned long long)x;
        if (k > vmax) k = vmax;
        return make_key<U>(k, 0);
    }
    case B200SORT_DISTR_STEP: {
        // m distinct values step*(1..m), step = MAX / (m + 1)  (distr.rs:82-99)
        unsigned long long m = (unsigned long long)g.param;
        if (m < 1) m = 1;
        const unsigned long long step = vmax / (m + 1);
        return make_key<U>(step * (1 + u % m), 0);
    }
    case B200SORT_DISTR_ASCENDING:
    case B200SORT_DISTR_DESCENDING: {
        // floor(i * 2^min(BITS,64) / N), N = param (total length of the stream)
        unsigned long long N = (unsigned long long)g.param;
        if (N < 1) N = 1;
        const int sh = BITS >= 64 ? 64 : BITS;
        unsigned __int128 v = ((unsigned __int128)i << sh) / N;
        unsigned long long k = (unsigned long long)v;
        if (k > vmax) k = vmax;
        if (g.distr == B200SORT_DISTR_DESCENDING) k = vmax - k;
        return make_key<U>(k, sizeof(U) > 8 ? k : 0);
    }
    case B200SORT_DISTR_FLOAT_TEST: {
        if (IS_FLOAT) {
            // Uniform[-1,1) (tests.rs:135,156) with the specials of tests.rs:139-143
            // (and more) at every `stride`-th index
            unsigned long long stride = (unsigned long long)g.param;
            if (stride < 1) stride = 1ull << 20;
            if (sizeof(U) == 4) {
                if (i % stride == 0) {
                    const uint32_t sp[12] = {0x00000000u, 0x80000000u, 0x7F800000u, 0xFF800000u,
                                             0x7FC00000u, 0xFFC00000u, 0x7F800001u, 0xFF800001u,
                                             0x00000001u, 0x80000001u, 0x7F7FFFFFu, 0xFF7FFFFFu};
                    return make_key<U>(sp[(i / stride) % 12], 0);
                }
                const float x = (float)(u >> 40) * (1.0f / 8388608.0f) - 1.0f;
                return make_key<U>(__float_as_uint(x), 0);
            } else {
                if (i % stride == 0) {
                    const unsigned long long sp[12] = {
                        0x0000000000000000ull, 0x8000000000000000ull, 0x7FF0000000000000ull, 0xFFF0000000000000ull,
                        0x7FF8000000000000ull, 0xFFF8000000000000ull, 0x7FF0000000000001ull, 0xFFF0000000000001ull,
                        0x0000000000000001ull, 0x8000000000000001ull, 0x7FEFFFFFFFFFFFFFull, 0xFFEFFFFFFFFFFFFFull};
                    return make_key<U>(sp[(i / stride) % 12], 0);
                }
                const double x = (double)(u >> 11) * (1.0 / 4503599627370496.0) - 1.0;
                return make_key<U>((unsigned long long)__double_as_longlong(x), 0);
            }
        }
        return make_key<U>(u, mix64(u));
    }
    default:  // UNIFORM: raw bits (Standard, distr.rs:42-46; rng.fill, tests.rs:28)
        return make_key<U>(u, mix64(u));
    }
}

template <typename U, bool IS_FLOAT>
__global__ void __launch_bounds__(256) generate_kernel(U* __restrict__ out, size_t n, GenParams g) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        out[i] = gen_one<U, IS_FLOAT>(g, g.index_base + i);
}

__device__ __forceinline__ unsigned long long warp_sum(unsigned long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
    return v;
}
__device__ __forceinline__ unsigned long long warp_xor(unsigned long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v ^= __shfl_xor_sync(0xFFFFFFFFu, v, o);
    return v;
}

template <typename U>
__device__ __forceinline__ unsigned long long key_hash(U k) { return mix64((unsigned long long)k); }
template <>
__device__ __forceinline__ unsigned long long key_hash<u128>(u128 k) { return mix64(k.lo ^ mix64(k.hi)); }

template <typename U>
__global__ void __launch_bounds__(256) fingerprint_kernel(const U* __restrict__ in, size_t n,
                                                          unsigned long long* __restrict__ out) {
    unsigned long long s1 = 0, x1 = 0, s2 = 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const U k = in[i];
        const unsigned long long h = key_hash<U>(k);
        s1 += h;
        x1 ^= h;
        s2 += low64<U>(k);
    }
    s1 = warp_sum(s1);
    x1 = warp_xor(x1);
    s2 = warp_sum(s2);
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&out[1], s1);
        atomicXor(&out[2], x1);
        atomicAdd(&out[3], s2);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = (unsigned long long)n;
}

template <typename U, int XF>
__global__ void __launch_bounds__(256) count_unsorted_kernel(const U* __restrict__ in, size_t n,
                                                             unsigned long long* __restrict__ out) {
    unsigned long long bad = 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i + 1 < n; i += stride) {
        const U a = to_ordered<U, XF>(in[i]);
        const U b = to_ordered<U, XF>(in[i + 1]);
        if (ordered_less<U>(b, a)) ++bad;
    }
    bad = warp_sum(bad);
    if ((threadIdx.x & 31) == 0 && bad) atomicAdd(out, bad);
}

unsigned grid_for(size_t n) {
    size_t b = (n + 255) / 256;
    if (b > 148 * 16) b = 148 * 16;
    if (b < 1) b = 1;
    return (unsigned)b;
}

int aux_fail(int code, const char* msg) {
    set_last_error(msg);
    return code;
}
int aux_cuda(cudaError_t e) {
    if (e == cudaSuccess) return B200SORT_OK;
    set_last_error(std::string("aux kernel: ") + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")");
    return B200SORT_ERR_CUDA;
}

#define AUX_FOR_KIND(kind, STMT_U, STMT_S, STMT_F)                                    \
    switch (kind) {                                                                   \
    case B200SORT_U8:   { using U = uint8_t; STMT_U; } break;                         \
    case B200SORT_U16:  { using U = uint16_t; STMT_U; } break;                        \
    case B200SORT_U32:  { using U = uint32_t; STMT_U; } break;                        \
    case B200SORT_U64:  { using U = unsigned long long; STMT_U; } break;              \
    case B200SORT_U128: { using U = u128; STMT_U; } break;                            \
    case B200SORT_I8:   { using U = uint8_t; STMT_S; } break;                         \
    case B200SORT_I16:  { using U = uint16_t; STMT_S; } break;                        \
    case B200SORT_I32:  { using U = uint32_t; STMT_S; } break;                        \
    case B200SORT_I64:  { using U = unsigned long long; STMT_S; } break;              \
    case B200SORT_I128: { using U = u128; STMT_S; } break;                            \
    case B200SORT_F32:  { using U = uint32_t; STMT_F; } break;                        \
    case B200SORT_F64:  { using U = unsigned long long; STMT_F; } break;              \
    default: return aux_fail(B200SORT_ERR_UNSUPPORTED_KIND, "unknown key kind");      \
    }

}  // namespace

extern "C" {

int b200sort_generate(void* d_keys, size_t n, int kind, int distr, double param, uint64_t seed,
                      uint64_t index_base, void* stream) {
    if (n == 0) return B200SORT_OK;
    if (!d_keys) return aux_fail(B200SORT_ERR_INVALID_ARG, "null device pointer");
    if (distr < B200SORT_DISTR_UNIFORM || distr > B200SORT_DISTR_FLOAT_TEST)
        return aux_fail(B200SORT_ERR_INVALID_ARG, "unknown distribution");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const GenParams g{distr, param, seed, index_base};
    const unsigned grid = grid_for(n);
    AUX_FOR_KIND(kind,
                 (generate_kernel<U, false><<<grid, 256, 0, st>>>(static_cast<U*>(d_keys), n, g)),
                 (generate_kernel<U, false><<<grid, 256, 0, st>>>(static_cast<U*>(d_keys), n, g)),
                 (generate_kernel<U, true><<<grid, 256, 0, st>>>(static_cast<U*>(d_keys), n, g)))
    return aux_cuda(cudaGetLastError());
}

int b200sort_fingerprint(const void* d_keys, size_t n, int kind, uint64_t* d_out, void* stream) {
    if (!d_out || (n && !d_keys)) return aux_fail(B200SORT_ERR_INVALID_ARG, "null device pointer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    cudaError_t e = cudaMemsetAsync(d_out, 0, 4 * sizeof(uint64_t), st);
    if (e != cudaSuccess) return aux_cuda(e);
    unsigned long long* out = reinterpret_cast<unsigned long long*>(d_out);
    const unsigned grid = grid_for(n);
    AUX_FOR_KIND(kind,
                 (fingerprint_kernel<U><<<grid, 256, 0, st>>>(static_cast<const U*>(d_keys), n, out)),
                 (fingerprint_kernel<U><<<grid, 256, 0, st>>>(static_cast<const U*>(d_keys), n, out)),
                 (fingerprint_kernel<U><<<grid, 256, 0, st>>>(static_cast<const U*>(d_keys), n, out)))
    return aux_cuda(cudaGetLastError());
}

int b200sort_count_unsorted(const void* d_keys, size_t n, int kind, uint64_t* d_out, void* stream) {
    if (!d_out || (n && !d_keys)) return aux_fail(B200SORT_ERR_INVALID_ARG, "null device pointer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    cudaError_t e = cudaMemsetAsync(d_out, 0, sizeof(uint64_t), st);
    if (e != cudaSuccess) return aux_cuda(e);
    if (n < 2) return B200SORT_OK;
    unsigned long long* out = reinterpret_cast<unsigned long long*>(d_out);
    const unsigned grid = grid_for(n);
    AUX_FOR_KIND(kind,
                 (count_unsorted_kernel<U, XF_NONE><<<grid, 256, 0, st>>>(static_cast<const U*>(d_keys), n, out)),
                 (count_unsorted_kernel<U, XF_SIGNED><<<grid, 256, 0, st>>>(static_cast<const U*>(d_keys), n, out)),
                 (count_unsorted_kernel<U, XF_FLOAT><<<grid, 256, 0, st>>>(static_cast<const U*>(d_keys), n, out)))
    return aux_cuda(cudaGetLastError());
}

}  // extern "C"
