// microbench3.cu -- shared-memory atomics on a per-warp 256-counter array when the
// lanes of a warp hit only K distinct counters (skewed digits): cost per warp
// instruction with and without a returned value.  Dev tool:
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/microbench3.cu -o /tmp/mb3 && /tmp/mb3
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

constexpr int THREADS = 512;
constexpr int ITERS = 2048;
__device__ __forceinline__ uint32_t lcg(uint32_t x) { return x * 1664525u + 1013904223u; }

template <int RET>
__global__ void __launch_bounds__(THREADS) bench(uint32_t* out, long long* cycles, int K) {
    __shared__ uint32_t sm[THREADS / 32 * 256];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < THREADS / 32 * 256; i += THREADS) sm[i] = 0;
    __syncthreads();
    uint32_t* my = sm + warp * 256;
    uint32_t x = lcg(blockIdx.x * THREADS + tid + 12345u), acc = 0;
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            x = lcg(x);
            const uint32_t d = (((x >> 24) % (unsigned)K) * 37u) & 255u;  // K distinct counters, distinct banks for K <= 32
            if (RET == 0) atomicAdd(&my[d], 1u);
            if (RET == 1) acc ^= atomicAdd(&my[d], 1u);
            if (RET == 2) {  // what a match-based ranker would do instead: group mask + leader atomic
                const unsigned m = __match_any_sync(0xFFFFFFFFu, d);
                const int leader = __ffs(m) - 1;
                uint32_t base = 0;
                if (lane == leader) base = atomicAdd(&my[d], (uint32_t)__popc(m));
                acc ^= __shfl_sync(0xFFFFFFFFu, base, leader) + __popc(m & ((1u << lane) - 1u));
            }
        }
    }
    const long long t1 = clock64();
    if (tid == 0) cycles[blockIdx.x] = t1 - t0;
    out[blockIdx.x * THREADS + tid] = acc + my[lane];
}

template <int RET>
void run(int K, uint32_t* out, long long* cyc, int bps) {
    const int blocks = 148 * bps;
    bench<RET><<<blocks, THREADS>>>(out, cyc, K);
    bench<RET><<<blocks, THREADS>>>(out, cyc, K);
    cudaDeviceSynchronize();
    long long h[148 * 4];
    cudaMemcpy(h, cyc, sizeof(long long) * blocks, cudaMemcpyDeviceToHost);
    double avg = 0;
    for (int i = 0; i < blocks; ++i) avg += (double)h[i];
    avg /= blocks;
    const double winst = (double)bps * (THREADS / 32) * ITERS * 8;
    const char* nm[] = {"atoms.add no-ret", "atoms.add ret", "match_any + leader atoms.add ret"};
    printf("%-34s K=%3d blocks/SM=%d : %8.2f cycles per warp-instr per SM\n", nm[RET], K, bps, avg / winst);
}

int main() {
    uint32_t* out;
    long long* cyc;
    cudaMalloc(&out, sizeof(uint32_t) * 148 * 4 * THREADS);
    cudaMalloc(&cyc, sizeof(long long) * 148 * 4);
    for (int K : {1, 2, 3, 4, 8, 16, 32, 64, 256}) {
        run<0>(K, out, cyc, 2);
        run<1>(K, out, cyc, 2);
        run<2>(K, out, cyc, 2);
    }
    cudaError_t e = cudaGetLastError();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
