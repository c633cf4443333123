// Yardstick for radix_pass_kernel (tools only, never linked into libpte.so): cub::DeviceRadixSort::SortKeys
// on the same shape the replay path sorts -- u64 records, 24 key bits at [38, 62) = three 8-bit passes --
// and the full 64 bits for reference.  Prints ms per call and per 8-bit pass.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_cub_sort tools/ubench_cub_sort.cu
// Run  : tools/ubench_cub_sort [records, default 1e9]
#include <cstdio>
#include <cstdlib>
#include <cub/cub.cuh>
typedef unsigned long long u64;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__global__ void fill(u64* a, u64 n) {
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        u64 z = i + 0x9E3779B97F4A7C15ull;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        a[i] = z ^ (z >> 31);
    }
}

int main(int argc, char** argv) {
    const u64 n = argc > 1 ? (u64)atof(argv[1]) : 1000000000ull;
    u64 *a, *b;
    CK(cudaMalloc(&a, n * 8)); CK(cudaMalloc(&b, n * 8));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const int ranges[3][2] = {{38, 62}, {0, 64}, {38, 46}};
    for (auto& r : ranges) {
        size_t tmp_bytes = 0;
        CK(cub::DeviceRadixSort::SortKeys(nullptr, tmp_bytes, a, b, (long long)n, r[0], r[1]));
        void* tmp = nullptr;
        CK(cudaMalloc(&tmp, tmp_bytes));
        float best = 1e9f;
        for (int rep = 0; rep < 4; rep++) {
            fill<<<1184, 256>>>(a, n);
            CK(cudaEventRecord(e0));
            CK(cub::DeviceRadixSort::SortKeys(tmp, tmp_bytes, a, b, (long long)n, r[0], r[1]));
            CK(cudaEventRecord(e1));
            CK(cudaDeviceSynchronize());
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            if (rep) best = fminf(best, ms);
        }
        const int passes = (r[1] - r[0] + 7) / 8;
        printf("cub::DeviceRadixSort::SortKeys u64 n=%llu bits [%d,%d): %.3f ms = %.3f ms per 8-bit pass (incl. its histogram kernel), temp %zu MB\n",
               n, r[0], r[1], best, best / passes, tmp_bytes >> 20);
        CK(cudaFree(tmp));
    }
    return 0;
}
