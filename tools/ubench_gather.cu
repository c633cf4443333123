// What does one random 16-byte dictionary probe cost on B200?  Random aligned loads over a table of a
// given size, four independent loads in flight per thread (the shape of intern_pack_kernel's probes).
// Prints useful GB/s and loads/s; run under `ncu --metrics dram__bytes_read.sum,lts__t_sectors_op_read.sum`
// to get DRAM bytes per load.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_gather tools/ubench_gather.cu
// Run  : tools/ubench_gather <table MB> [fetch-granularity limit]
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
typedef unsigned long long u64;
typedef unsigned int u32;

__device__ __forceinline__ u64 mix64(u64 z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

enum { PLAIN16, CG16, NC16, EVICT_FIRST16, EVICT_LAST16, NOALLOC16, PF64_16, PF128_16, PLAIN8, PAIR32, CV16, CAS128, CAS64, RELAXED16, N_MODES };
const char* kNames[N_MODES] = {"ld.global.v2.u64", "ld.global.cg.v2.u64", "ld.global.nc.v2.u64", "cache_hint evict_first", "cache_hint evict_last",
                               "L1::no_allocate", "ld.global.L2::64B", "ld.global.L2::128B", "ld.global.u64 (8 B)", "2 x 16 B in one sector", "ld.volatile (cv)",
                               "atom.cas.b128 (never matches)", "atom.cas.b64 (never matches)", "ld.relaxed.gpu.v2.u64"};

template <int MODE>
__device__ __forceinline__ u64 probe(const ulonglong2* t, u64 slot, u64 pol) {
    const ulonglong2* p = t + slot;
    (void)p;
    ulonglong2 v = make_ulonglong2(0, 0);
    if (MODE == PLAIN16) asm volatile("ld.global.v2.u64 {%0,%1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p));
    if (MODE == CG16) asm volatile("ld.global.cg.v2.u64 {%0,%1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p));
    if (MODE == CV16) asm volatile("ld.volatile.global.v2.u64 {%0,%1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p));
    if (MODE == NC16) asm volatile("ld.global.nc.v2.u64 {%0,%1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p));
    if (MODE == EVICT_FIRST16 || MODE == EVICT_LAST16)
        asm volatile("ld.global.L2::cache_hint.v2.u64 {%0,%1}, [%2], %3;" : "=l"(v.x), "=l"(v.y) : "l"(p), "l"(pol));
    if (MODE == NOALLOC16) asm volatile("ld.global.L1::no_allocate.v2.u64 {%0,%1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p));
    if (MODE == PF64_16) asm volatile("ld.global.L2::64B.v2.u64 {%0,%1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p));
    if (MODE == PF128_16) asm volatile("ld.global.L2::128B.v2.u64 {%0,%1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p));
    if (MODE == CAS128) {         // served by the slice that owns the line: does the other die keep a copy of what an atomic read?
        asm volatile("{ .reg .b128 c, n, o; mov.b128 c, {%3, %3}; mov.b128 n, {%4, %4}; atom.global.cas.b128 o, [%2], c, n; mov.b128 {%0, %1}, o; }"
                     : "=l"(v.x), "=l"(v.y) : "l"(p), "l"(0ull), "l"(5ull) : "memory");
    }
    if (MODE == CAS64) asm volatile("atom.global.cas.b64 %0, [%1], %2, %3;" : "=l"(v.x) : "l"(p), "l"(0ull), "l"(5ull) : "memory");
    if (MODE == RELAXED16) asm volatile("ld.relaxed.gpu.global.v2.u64 {%0,%1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p) : "memory");
    if (MODE == PLAIN8) asm volatile("ld.global.u64 %0, [%1];" : "=l"(v.x) : "l"(p));
    if (MODE == PAIR32) {
        const ulonglong2* q = t + (slot & ~1ull);
        ulonglong2 w;
        asm volatile("ld.global.v2.u64 {%0,%1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(q));
        asm volatile("ld.global.v2.u64 {%0,%1}, [%2];" : "=l"(w.x), "=l"(w.y) : "l"(q + 1));
        v.x ^= w.x; v.y ^= w.y;
    }
    return v.x ^ v.y;
}

template <int MODE, int ILP = 4>
__global__ void __launch_bounds__(256) gather(const ulonglong2* __restrict__ t, u64 slots, u64 per_thread, u64 seed, u64* out) {
    u64 pol = 0;
    if (MODE == EVICT_FIRST16) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    if (MODE == EVICT_LAST16) asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    const u64 gid = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    u64 acc = 0;
    for (u64 i = 0; i < per_thread; i += ILP) {
        u64 s[ILP];
#pragma unroll
        for (int k = 0; k < ILP; k++) s[k] = __umul64hi(mix64(seed + (gid * per_thread + i + k) * 0x9E3779B97F4A7C15ull), slots);
        u64 v[ILP];
#pragma unroll
        for (int k = 0; k < ILP; k++) v[k] = probe<MODE>(t, s[k], pol);
#pragma unroll
        for (int k = 0; k < ILP; k++) acc += v[k];
    }
    if (acc == 0x1234567ull) out[0] = acc;      // keeps the loads alive
}

template <int MODE, int ILP = 4>
void run(const ulonglong2* t, u64 slots, u64* out, int ctas_per_sm = 8) {
    const int blocks = 148 * ctas_per_sm;
    const u64 per_thread = 1024;
    const u64 loads = (u64)blocks * 256 * per_thread;
    gather<MODE, ILP><<<blocks, 256>>>(t, slots, per_thread, 1, out);
    cudaDeviceSynchronize();
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a);
    gather<MODE, ILP><<<blocks, 256>>>(t, slots, per_thread, 2, out);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    const double bytes = MODE == PLAIN8 ? 8 : (MODE == PAIR32 ? 32 : 16);
    printf("  %-26s ilp %2d x %d CTAs/SM %8.3f ms  %7.2f G loads/s  %7.1f GB/s useful\n", kNames[MODE], ILP, ctas_per_sm, ms, loads / ms * 1e-6, loads * bytes / ms * 1e-6);
}

int main(int argc, char** argv) {
    const u64 mb = argc > 1 ? strtoull(argv[1], nullptr, 10) : 256;
    if (argc > 2) {
        cudaError_t e = cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)atoi(argv[2]));
        size_t got = 0;
        cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity);
        printf("fetch-granularity limit: asked %s, set -> %s, now %zu\n", argv[2], cudaGetErrorString(e), got);
    } else {
        size_t got = 0;
        cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity);
        printf("fetch-granularity limit (default): %zu\n", got);
    }
    u64 slots = (mb << 20) / 16;             // any size: slot = hash % slots when not a power of two
    ulonglong2* t;
    u64* out;
    if (cudaMalloc(&t, slots * 16) != cudaSuccess || cudaMalloc(&out, 8) != cudaSuccess) { printf("alloc failed\n"); return 1; }
    cudaMemset(t, 1, slots * 16);
    printf("table %llu MB (%llu slots of 16 B), %d x 256 threads, 4 loads in flight per thread\n", slots * 16 >> 20, slots, 148 * 8);
    run<PLAIN16>(t, slots, out);
    if (argc > 3 && argv[3][0] == 'i') {                       // loads in flight: per thread x CTAs per SM
        run<PLAIN16, 1>(t, slots, out); run<PLAIN16, 2>(t, slots, out); run<PLAIN16, 8>(t, slots, out); run<PLAIN16, 16>(t, slots, out);
        run<PLAIN16, 4>(t, slots, out, 2); run<PLAIN16, 4>(t, slots, out, 4); run<PLAIN16, 4>(t, slots, out, 5);
        run<PLAIN16, 8>(t, slots, out, 4); run<PLAIN16, 16>(t, slots, out, 2);
    }
    if (argc > 3) { cudaDeviceSynchronize(); return 0; }      // size sweeps: the plain load only
    run<CG16>(t, slots, out);
    run<CV16>(t, slots, out);
    run<NC16>(t, slots, out);
    run<NOALLOC16>(t, slots, out);
    run<EVICT_FIRST16>(t, slots, out);
    run<EVICT_LAST16>(t, slots, out);
    run<PF64_16>(t, slots, out);
    run<PF128_16>(t, slots, out);
    run<PLAIN8>(t, slots, out);
    run<PAIR32>(t, slots, out);
    run<CAS128>(t, slots, out);
    run<CAS64>(t, slots, out);
    run<RELAXED16>(t, slots, out);
    cudaError_t e = cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(e));
    return e == cudaSuccess ? 0 : 1;
}
