// microbench2.cu -- (a) do shared-memory atomics on B200 resolve same-address lanes of
// one warp instruction in ascending lane order?  (b) throughput of data-dependent
// same-address shared atomics (skewed digits).  Dev tool.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#define FULL 0xFFFFFFFFu
constexpr int THREADS = 512;

__device__ __forceinline__ uint32_t lcg(uint32_t x) { return x * 1664525u + 1013904223u; }

// Every warp does `rounds` rounds of old = atomicAdd(&cnt[d], 1) on warp-private counters and
// checks old against the lane-ordered rank computed with ballots.
__global__ void __launch_bounds__(THREADS) order_test(unsigned long long* viol, unsigned long long* dups, int rounds, int kmask) {
    __shared__ uint32_t cnt[THREADS / 32 * 256];
    __shared__ uint32_t ref[THREADS / 32 * 256];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < THREADS / 32 * 256; i += THREADS) { cnt[i] = 0; ref[i] = 0; }
    __syncthreads();
    uint32_t* my = cnt + warp * 256;
    uint32_t* myref = ref + warp * 256;
    uint32_t x = lcg(blockIdx.x * THREADS + tid + 777u);
    unsigned long long v = 0, dd = 0;
    for (int r = 0; r < rounds; ++r) {
        x = lcg(x);
        const uint32_t d = (x >> 24) & kmask;
        const uint32_t old = atomicAdd(&my[d], 1u);
        unsigned m = FULL;
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            const unsigned bal = __ballot_sync(FULL, (d >> b) & 1);
            m &= ((d >> b) & 1) ? bal : ~bal;
        }
        const int leader = __ffs(m) - 1;
        uint32_t base = 0;
        if (lane == leader) { base = myref[d]; myref[d] = base + __popc(m); }
        base = __shfl_sync(FULL, base, leader);
        const uint32_t want = base + __popc(m & ((1u << lane) - 1u));
        if (old != want) ++v;
        if (__popc(m) > 1) ++dd;
        __syncwarp();
        if ((r & 127) == 127) {  // keep counters small
            __syncwarp();
            for (int i = lane; i < 256; i += 32) { my[i] = 0; myref[i] = 0; }
            __syncwarp();
        }
    }
    atomicAdd(viol, v);
    atomicAdd(dups, dd);
}

template <bool RET>
__global__ void __launch_bounds__(THREADS) skew_atoms(uint32_t* out, long long* cycles, int kmask, int iters) {
    __shared__ uint32_t sm[THREADS / 32 * 256];
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < THREADS / 32 * 256; i += THREADS) sm[i] = 0;
    __syncthreads();
    uint32_t* my = sm + warp * 256;
    uint32_t x = lcg(blockIdx.x * THREADS + tid + 99u), acc = 0;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            x = lcg(x);
            const uint32_t d = (x >> 24) & kmask;
            if (RET) acc ^= atomicAdd(&my[d], 1u); else atomicAdd(&my[d], 1u);
        }
    }
    const long long t1 = clock64();
    if (tid == 0) cycles[blockIdx.x] = t1 - t0;
    out[blockIdx.x * THREADS + tid] = acc + my[tid & 255];
}

int main() {
    unsigned long long *d, h[2];
    cudaMalloc(&d, 16);
    for (int kmask : {0xFF, 0x0F, 0x03, 0x01, 0x00}) {
        cudaMemset(d, 0, 16);
        order_test<<<148 * 4, THREADS>>>(d, d + 1, 20000, kmask);
        cudaDeviceSynchronize();
        cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        printf("order test kmask=0x%02x: lane-rounds=%llu  in-dup-group=%llu  order violations=%llu\n", kmask,
               (unsigned long long)148 * 4 * THREADS * 20000ull, h[1], h[0]);
    }
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, 4 * 148 * 4 * THREADS);
    cudaMalloc(&cyc, 8 * 148 * 4);
    const int iters = 2048, blocks = 148 * 4;
    for (int ret = 0; ret < 2; ++ret)
        for (int kmask : {0xFF, 0x3F, 0x0F, 0x03, 0x01, 0x00}) {
            for (int rep = 0; rep < 2; ++rep) {
                if (ret) skew_atoms<true><<<blocks, THREADS>>>(out, cyc, kmask, iters);
                else skew_atoms<false><<<blocks, THREADS>>>(out, cyc, kmask, iters);
            }
            cudaDeviceSynchronize();
            long long hc[148 * 4];
            cudaMemcpy(hc, cyc, 8 * blocks, cudaMemcpyDeviceToHost);
            double avg = 0;
            for (int i = 0; i < blocks; ++i) avg += hc[i];
            avg /= blocks;
            const double winst = 4.0 * (THREADS / 32) * iters * 8;
            printf("atoms.add %s, digits over %3d values: %7.2f cycles per warp-instr per SM\n", ret ? "ret  " : "noret", kmask + 1, avg / winst);
        }
    printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
