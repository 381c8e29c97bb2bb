// microbench.cu -- per-SM throughput of the warp/shared-memory primitives a radix
// ranking step can be built from, measured on the box (dev tool, not product).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o microbench microbench.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>

#define FULL 0xFFFFFFFFu
constexpr int THREADS = 512;
constexpr int ITERS = 2048;

__device__ __forceinline__ uint32_t lcg(uint32_t x) { return x * 1664525u + 1013904223u; }

enum Op { OP_BALLOT, OP_MATCH_RANDOM8, OP_MATCH_K, OP_SHFL, OP_ATOMS_ADD, OP_ATOMS_ADD_RET, OP_ATOMS_OR,
          OP_LDS_RANDOM, OP_STS_RANDOM, OP_LDS_LINEAR, OP_REDUX_OR, OP_POPC, OP_ATOMS_SAME, OP_MATCH_ALL,
          OP_LDS64_RANDOM, OP_STS64_RANDOM, OP_RMW_RANDOM, OP_COUNT };
const char* names[] = {"ballot", "match_any(random 8-bit)", "match_any(K distinct)", "shfl_idx", "atoms.add random (no ret)",
                       "atoms.add random (ret)", "atoms.or random (no ret)", "lds random bank", "sts random bank",
                       "lds linear", "redux.or", "popc+ffs", "atoms.add same addr", "match_all", "lds.64 random",
                       "sts.64 random", "lds+sts rmw random"};

template <int OP>
__global__ void __launch_bounds__(THREADS) bench(uint32_t* out, long long* cycles, int K) {
    __shared__ uint32_t sm[THREADS / 32 * 256 * 2];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < THREADS / 32 * 256 * 2; i += THREADS) sm[i] = i;
    __syncthreads();
    uint32_t* my = sm + warp * 256;
    unsigned long long* my64 = reinterpret_cast<unsigned long long*>(sm) + warp * 256;
    uint32_t x = lcg(blockIdx.x * THREADS + tid + 12345u), acc = 0;
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            x = lcg(x);
            const uint32_t d = x >> 24;
            if (OP == OP_BALLOT) acc ^= __ballot_sync(FULL, (d >> u) & 1);
            if (OP == OP_MATCH_RANDOM8) acc ^= __match_any_sync(FULL, d);
            if (OP == OP_MATCH_K) acc ^= __match_any_sync(FULL, (d % (unsigned)K));
            if (OP == OP_SHFL) acc ^= __shfl_sync(FULL, x, d & 31);
            if (OP == OP_ATOMS_ADD) atomicAdd(&my[d], 1u);
            if (OP == OP_ATOMS_ADD_RET) acc ^= atomicAdd(&my[d], 1u);
            if (OP == OP_ATOMS_OR) atomicOr(&my[d], 1u << lane);
            if (OP == OP_LDS_RANDOM) acc ^= my[d];
            if (OP == OP_STS_RANDOM) my[d] = x;
            if (OP == OP_LDS_LINEAR) acc ^= my[(lane + u * 32) & 255];
            if (OP == OP_REDUX_OR) acc ^= __reduce_or_sync(FULL, d);
            if (OP == OP_POPC) acc += __popc(x) + __ffs(x ^ acc);
            if (OP == OP_ATOMS_SAME) atomicAdd(&my[u], 1u);
            if (OP == OP_MATCH_ALL) { int p; acc ^= __match_all_sync(FULL, d & 1, &p); acc += p; }
            if (OP == OP_LDS64_RANDOM) acc ^= (uint32_t)my64[d];
            if (OP == OP_STS64_RANDOM) my64[d] = x;
            if (OP == OP_RMW_RANDOM) { my[d] = my[d] + 1; __syncwarp(); }
        }
    }
    const long long t1 = clock64();
    if (tid == 0) cycles[blockIdx.x] = t1 - t0;
    out[blockIdx.x * THREADS + tid] = acc + my[lane];
}

template <int OP>
void run(int K, uint32_t* out, long long* cyc, int blocks_per_sm) {
    const int sms = 148, blocks = sms * blocks_per_sm;
    bench<OP><<<blocks, THREADS>>>(out, cyc, K);
    bench<OP><<<blocks, THREADS>>>(out, cyc, K);
    cudaDeviceSynchronize();
    long long h[148 * 4];
    cudaMemcpy(h, cyc, sizeof(long long) * blocks, cudaMemcpyDeviceToHost);
    double avg = 0;
    for (int i = 0; i < blocks; ++i) avg += (double)h[i];
    avg /= blocks;
    // warp-instructions of the op issued per SM while the block ran
    const double winst = (double)blocks_per_sm * (THREADS / 32) * ITERS * 8;
    printf("%-30s K=%2d blocks/SM=%d : %8.2f cycles per warp-instr per SM  (%6.3f warp-instr/cycle/SM)\n", names[OP], K,
           blocks_per_sm, avg / winst, winst / avg);
}

int main() {
    uint32_t* out;
    long long* cyc;
    cudaMalloc(&out, sizeof(uint32_t) * 148 * 4 * THREADS);
    cudaMalloc(&cyc, sizeof(long long) * 148 * 4);
    for (int bps = 2; bps <= 4; bps += 2) {
        run<OP_BALLOT>(0, out, cyc, bps);
        run<OP_MATCH_RANDOM8>(0, out, cyc, bps);
        for (int K : {1, 2, 4, 8, 16, 32}) run<OP_MATCH_K>(K, out, cyc, bps);
        run<OP_MATCH_ALL>(0, out, cyc, bps);
        run<OP_SHFL>(0, out, cyc, bps);
        run<OP_REDUX_OR>(0, out, cyc, bps);
        run<OP_POPC>(0, out, cyc, bps);
        run<OP_ATOMS_ADD>(0, out, cyc, bps);
        run<OP_ATOMS_ADD_RET>(0, out, cyc, bps);
        run<OP_ATOMS_OR>(0, out, cyc, bps);
        run<OP_ATOMS_SAME>(0, out, cyc, bps);
        run<OP_LDS_RANDOM>(0, out, cyc, bps);
        run<OP_STS_RANDOM>(0, out, cyc, bps);
        run<OP_RMW_RANDOM>(0, out, cyc, bps);
        run<OP_LDS64_RANDOM>(0, out, cyc, bps);
        run<OP_STS64_RANDOM>(0, out, cyc, bps);
        run<OP_LDS_LINEAR>(0, out, cyc, bps);
    }
    cudaError_t e = cudaGetLastError();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
