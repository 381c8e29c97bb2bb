// cub_compare.cu -- same-box comparator (bench-only; never linked into the product):
// CUB DeviceRadixSort::SortKeys (onesweep, sm_100 tuning) on uniform keys.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o cub_compare cub_compare.cu
#include <cub/cub.cuh>
#include <cstdio>
#include <cstdint>
#include <cstdlib>

__global__ void fill(unsigned long long* p, size_t n, unsigned long long seed) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, s = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += s) {
        unsigned long long z = seed ^ (i * 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        p[i] = z ^ (z >> 31);
    }
}

template <typename K>
void run(size_t n, const char* name) {
    K *in, *out;
    cudaMalloc(&in, n * sizeof(K));
    cudaMalloc(&out, n * sizeof(K));
    fill<<<148 * 8, 256>>>((unsigned long long*)in, n * sizeof(K) / 8, 1234);
    size_t tb = 0;
    cub::DeviceRadixSort::SortKeys(nullptr, tb, in, out, n);
    void* tmp;
    cudaMalloc(&tmp, tb);
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    float best = 1e30f;
    for (int r = 0; r < 6; ++r) {
        cudaEventRecord(a);
        cub::DeviceRadixSort::SortKeys(tmp, tb, in, out, n);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms;
        cudaEventElapsedTime(&ms, a, b);
        if (r > 0 && ms < best) best = ms;
    }
    printf("{\"impl\": \"cub\", \"kind\": \"%s\", \"n\": %zu, \"total_ms\": %.4f, \"gkeys_s\": %.3f, \"tmp_bytes\": %zu}\n", name, n,
           best, n / best / 1e6, tb);
    cudaFree(in); cudaFree(out); cudaFree(tmp);
}

int main(int argc, char** argv) {
    size_t n = argc > 1 ? (size_t)atof(argv[1]) : (size_t)1 << 28;
    run<uint32_t>(n, "u32");
    run<unsigned long long>(n, "u64");
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) printf("error: %s\n", cudaGetErrorString(e));
    return 0;
}
