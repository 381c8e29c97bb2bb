// microbench4.cu -- can the write-out of a K3 tile go through the TMA engine?  A CTA
// issues one cp.async.bulk (shared -> global) per digit run -- 256 copies of RUN
// bytes to scattered 16-byte aligned global addresses per "tile" -- and waits for
// the reads of shared memory to finish, like a one-tile kernel would before it
// exits.  Reports tiles/us/SM and GB/s for 2 CTAs per SM.  Dev tool:
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/microbench4.cu -o tools/bin/mb4 && tools/bin/mb4
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

constexpr int THREADS = 512;
__device__ __forceinline__ uint32_t lcg(uint32_t x) { return x * 1664525u + 1013904223u; }
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// MODE 0: 256 threads issue one bulk copy each.  MODE 1: the same bytes by plain
// per-thread 8-byte stores (what the kernel does today), for comparison.
template <int MODE>
__global__ void __launch_bounds__(THREADS, 2) bench(unsigned char* out, size_t out_bytes, int run_bytes, int tiles) {
    extern __shared__ __align__(128) unsigned char tile[];
    const int tid = threadIdx.x;
    for (int i = tid; i < 256 * run_bytes / 8; i += THREADS) reinterpret_cast<unsigned long long*>(tile)[i] = i;
    __syncthreads();
    uint32_t x = lcg(blockIdx.x * 977u + tid * 31u + 7u);
    const size_t slots = out_bytes / 4096;  // scattered 4 KB-aligned destinations, run lands at a 16-byte offset
    for (int t = 0; t < tiles; ++t) {
        if (MODE == 0) {
            if (tid < 256) {
                x = lcg(x);
                unsigned char* dst = out + (size_t)(x % slots) * 4096 + ((x >> 20) & 15) * 16;
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                             ::"l"(dst), "r"(smem_u32(tile + tid * run_bytes)), "r"(run_bytes) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            }
            __syncthreads();
        } else {
            // warp w copies runs w, w+16, ...: 32 lanes x 8 bytes per instruction
            const int lane = tid & 31, warp = tid >> 5;
            for (int r = warp; r < 256; r += THREADS / 32) {
                uint32_t y = lcg(blockIdx.x * 977u + r * 31u + t * 131u);
                unsigned char* dst = out + (size_t)(y % slots) * 4096 + ((y >> 20) & 15) * 16;
                for (int b = lane * 8; b < run_bytes; b += 256)
                    *reinterpret_cast<unsigned long long*>(dst + b) = *reinterpret_cast<unsigned long long*>(tile + r * run_bytes + b);
            }
            __syncthreads();
        }
    }
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const size_t out_bytes = 4ull << 30;
    unsigned char* out;
    cudaMalloc(&out, out_bytes);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int tiles = 400;
    for (int mode = 0; mode < 2; ++mode)
        for (int run : {64, 128, 256, 512}) {
            const size_t smem = 256 * (size_t)run;
            auto k = mode == 0 ? bench<0> : bench<1>;
            cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            k<<<sms * 2, THREADS, smem>>>(out, out_bytes, run, 8);
            cudaEventRecord(e0);
            k<<<sms * 2, THREADS, smem>>>(out, out_bytes, run, tiles);
            cudaEventRecord(e1);
            cudaError_t err = cudaDeviceSynchronize();
            float ms = 0;
            cudaEventElapsedTime(&ms, e0, e1);
            const double bytes = (double)sms * 2 * tiles * 256 * run;
            printf("%s run %4d B: %8.3f ms  %7.1f GB/s  %6.2f us per tile per CTA  %6.1f cycles per copy per SM (at 1.965 GHz)  %s\n",
                   mode == 0 ? "cp.async.bulk" : "thread stores", run, ms, bytes / ms / 1e6, ms * 1e3 / tiles,
                   ms * 1e-3 * 1.965e9 / (2.0 * tiles * 256), cudaGetErrorString(err));
        }
    return 0;
}
