// How wide can ONE scatter pass of 24-byte event records be on B200 before the write side falls apart?
// Events (uid_hi, uid_lo, phase, type) -> records (uid_hi, uid_lo, pos<<8|type<<6|phase) grouped by the top
// BITS bits of mix64(uid_lo).  Ranking inside a tile is by shared-memory atomicAdd (not stable: this tool
// measures the memory side only), the tile is staged in bin order in shared memory and written out word by
// word, so a (tile, bin) run leaves as one contiguous piece.  Global placement: MODE 0 takes each run's
// place with an atomicAdd on a per-bin cursor (tiles in flight fill neighbouring places, the pattern a
// look-back chain gives); MODE 1 gives every CTA one contiguous chunk of the input and offsets from a
// histogram pass (no chain, but CTAs x bins open write fronts).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_scatter tools/ubench_scatter.cu
// Run  : tools/ubench_scatter [events, default 2^28]
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
typedef unsigned long long u64;
typedef unsigned int u32;
typedef unsigned short u16;
typedef unsigned char u8;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__host__ __device__ __forceinline__ u64 mix64(u64 z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__global__ void gen_kernel(u64* hi, u64* lo, u8* ph, u8* ty, u64 n, u64 pods) {
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        u64 r = mix64(i + 0x9E3779B97F4A7C15ull);
        u64 pod = __umul64hi(r, pods);
        hi[i] = mix64(pod ^ 0x6B38735F77617463ull);
        lo[i] = mix64(~pod ^ 0x5F62323030706F64ull);
        ph[i] = (u8)((r >> 3) % 5);
        ty[i] = (u8)((r >> 11) & 1);
    }
}

template <int BITS>
__device__ __forceinline__ u32 bin_of(u64 lo) { return (u32)(mix64(lo) >> (64 - BITS)); }

template <int BITS>
__global__ void __launch_bounds__(512) hist_kernel(const u64* __restrict__ lo, u64 n, u64 chunk, u32* __restrict__ hist /*[grid][NB]*/) {
    constexpr int NB = 1 << BITS;
    extern __shared__ u32 sh[];
    for (int k = threadIdx.x; k < NB; k += blockDim.x) sh[k] = 0;
    __syncthreads();
    const u64 b = (u64)blockIdx.x * chunk, e = min(n, b + chunk);
    for (u64 i = b + threadIdx.x * 2; i < e; i += blockDim.x * 2) {
        const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(lo + i);
        atomicAdd(&sh[bin_of<BITS>(v.x)], 1u);
        if (i + 1 < e) atomicAdd(&sh[bin_of<BITS>(v.y)], 1u);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < NB; k += blockDim.x) hist[(u64)blockIdx.x * NB + k] = sh[k];
}

// bin-major, chunk-minor exclusive scan of hist[grid][NB] -> off[grid][NB]; also bin starts.  One block, slow but tiny.
__global__ void scan_kernel(const u32* hist, u32* off, u32* bin_start, int grid, int NB) {
    __shared__ u32 s_tot[2048];
    for (int b = threadIdx.x; b < NB; b += blockDim.x) {
        u32 t = 0;
        for (int c = 0; c < grid; c++) t += hist[(u64)c * NB + b];
        s_tot[b] = t;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        u32 run = 0;
        for (int b = 0; b < NB; b++) { u32 t = s_tot[b]; s_tot[b] = run; bin_start[b] = run; run += t; }
    }
    __syncthreads();
    for (int b = threadIdx.x; b < NB; b += blockDim.x) {
        u32 run = s_tot[b];
        for (int c = 0; c < grid; c++) { u32 t = hist[(u64)c * NB + b]; off[(u64)c * NB + b] = run; run += t; }
    }
}

template <int THREADS>
__device__ __forceinline__ u32 block_excl_scan(u32 v, u32* part /*THREADS/32*/, u32* total) {
    const u32 lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    u32 inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { u32 t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) part[w] = inc;
    __syncthreads();
    u32 base = 0, all = 0;
    for (int k = 0; k < THREADS / 32; k++) { u32 t = part[k]; if (k < (int)w) base += t; all += t; }
    if (total) *total = all;
    __syncthreads();
    return base + inc - v;
}

template <int BITS, int THREADS, int ITEMS, int MODE>
__global__ void __launch_bounds__(THREADS) scatter_kernel(const u64* __restrict__ hi, const u64* __restrict__ lo, const u8* __restrict__ ph,
                                                          const u8* __restrict__ ty, u64 n, u64 chunk, u64* __restrict__ out,
                                                          u32* __restrict__ gcursor, const u32* __restrict__ off, u32* __restrict__ ticket) {
    constexpr int NB = 1 << BITS, TILE = THREADS * ITEMS;
    constexpr int PER = NB / THREADS > 0 ? NB / THREADS : 1;
    extern __shared__ __align__(16) unsigned char raw[];
    u64* stage = reinterpret_cast<u64*>(raw);                    // TILE * 3
    u32* cnt = reinterpret_cast<u32*>(stage + TILE * 3);         // NB: counts, then exclusive starts inside the tile
    u32* gbase = cnt + NB;                                       // NB: global record index of the tile's run minus its start in the tile
    u32* run_off = gbase + NB;                                   // NB: MODE 1 running offsets
    u16* sbin = reinterpret_cast<u16*>(run_off + NB);            // TILE
    __shared__ u32 part[THREADS / 32];
    __shared__ u32 s_tile;
    const u32 tid = threadIdx.x;
    u64 c_beg = 0, c_end = n;
    if (MODE == 1) {
        c_beg = (u64)blockIdx.x * chunk; c_end = min(n, c_beg + chunk);
        for (int k = tid; k < NB; k += THREADS) run_off[k] = off[(u64)blockIdx.x * NB + k];
    }
    u64 tile_base = MODE == 1 ? c_beg : 0;
    for (;;) {
        if (MODE == 0) {
            if (tid == 0) s_tile = atomicAdd(ticket, 1u);
            __syncthreads();
            tile_base = (u64)s_tile * TILE;
        }
        if (tile_base >= c_end) break;
        const u32 valid = (u32)min((u64)TILE, c_end - tile_base);
        for (int k = tid; k < NB; k += THREADS) cnt[k] = 0;
        __syncthreads();
        u64 rhi[ITEMS], rlo[ITEMS], meta[ITEMS];
        u32 bin[ITEMS], rk[ITEMS];
#pragma unroll
        for (int k = 0; k < ITEMS; k++) {
            const u32 j = k * THREADS + tid;
            const u64 i = tile_base + j;
            if (j < valid) { rlo[k] = __ldcs(lo + i); rhi[k] = __ldcs(hi + i); meta[k] = (i << 8) | ((u64)__ldcs(ty + i) << 6) | __ldcs(ph + i); }
        }
#pragma unroll
        for (int k = 0; k < ITEMS; k++) {
            const u32 j = k * THREADS + tid;
            if (j < valid) { bin[k] = bin_of<BITS>(rlo[k]); rk[k] = atomicAdd(&cnt[bin[k]], 1u); }
        }
        __syncthreads();
        // exclusive scan of the bin counts
        u32 loc[PER], sum = 0;
        if (tid * PER < NB) {
#pragma unroll
            for (int q = 0; q < PER; q++) { loc[q] = cnt[tid * PER + q]; sum += loc[q]; }
        }
        u32 ex = block_excl_scan<THREADS>(sum, part, nullptr);
        if (tid * PER < NB) {
#pragma unroll
            for (int q = 0; q < PER; q++) {
                const u32 b = tid * PER + q, c = loc[q];
                cnt[b] = ex;
                if (c) {
                    u32 g;
                    if (MODE == 0) g = atomicAdd(&gcursor[b], c);
                    else { g = run_off[b]; run_off[b] = g + c; }
                    gbase[b] = g - ex;
                }
                ex += c;
            }
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < ITEMS; k++) {
            const u32 j = k * THREADS + tid;
            if (j < valid) {
                const u32 p = cnt[bin[k]] + rk[k];
                stage[p * 3] = rhi[k]; stage[p * 3 + 1] = rlo[k]; stage[p * 3 + 2] = meta[k];
                sbin[p] = (u16)bin[k];
            }
        }
        __syncthreads();
        for (u32 w = tid; w < valid * 3; w += THREADS) {
            const u32 rec = w / 3;
            out[(u64)(gbase[sbin[rec]] + rec) * 3 + (w - rec * 3)] = stage[w];
        }
        __syncthreads();
        if (MODE == 1) tile_base += TILE;
    }
}

template <int BITS>
__global__ void check_kernel(const u64* out, u64 n, const u32* bin_start, u32* bad) {
    constexpr int NB = 1 << BITS;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        const u32 b = bin_of<BITS>(out[i * 3 + 1]);
        const u64 lo_ = bin_start[b], hi_ = b + 1 < NB ? bin_start[b + 1] : n;
        if (i < lo_ || i >= hi_) atomicAdd(bad, 1u);
    }
}

struct Bufs { u64 *hi, *lo, *out; u8 *ph, *ty; u32 *hist, *off, *bin_start, *gcursor, *ticket, *bad; };

template <int BITS, int THREADS, int ITEMS, int MODE>
void run(const Bufs& B, u64 n, int sms) {
    constexpr int NB = 1 << BITS, TILE = THREADS * ITEMS;
    const size_t smem = (size_t)TILE * 24 + 3 * NB * 4 + TILE * 2;
    auto kern = scatter_kernel<BITS, THREADS, ITEMS, MODE>;
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) { cudaGetLastError(); printf("bits %d T %d I %d: smem %zu too large\n", BITS, THREADS, ITEMS, smem); return; }
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, THREADS, smem));
    if (occ < 1) { printf("does not fit\n"); return; }
    const int grid = sms * occ;
    const u64 chunk = ((n + grid - 1) / grid + 1) & ~1ull;
    cudaEvent_t e0, e1, e2;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1)); CK(cudaEventCreate(&e2));
    float best_h = 1e9f, best_s = 1e9f;
    for (int rep = 0; rep < 4; rep++) {
        CK(cudaMemset(B.ticket, 0, 4));
        CK(cudaEventRecord(e0));
        hist_kernel<BITS><<<grid, 512, NB * 4>>>(B.lo, n, chunk, B.hist);
        scan_kernel<<<1, 1024>>>(B.hist, B.off, B.bin_start, grid, NB);
        CK(cudaMemcpyAsync(B.gcursor, B.bin_start, NB * 4, cudaMemcpyDeviceToDevice));
        CK(cudaEventRecord(e1));
        kern<<<grid, THREADS, smem>>>(B.hi, B.lo, B.ph, B.ty, n, chunk, B.out, B.gcursor, B.off, B.ticket);
        CK(cudaEventRecord(e2));
        CK(cudaDeviceSynchronize());
        float h, s;
        cudaEventElapsedTime(&h, e0, e1); cudaEventElapsedTime(&s, e1, e2);
        if (rep) { best_h = fminf(best_h, h); best_s = fminf(best_s, s); }
    }
    CK(cudaMemset(B.bad, 0, 4));
    check_kernel<BITS><<<sms * 8, 256>>>(B.out, n, B.bin_start, B.bad);
    u32 bad = 0;
    CK(cudaMemcpy(&bad, B.bad, 4, cudaMemcpyDeviceToHost));
    const double gb = (double)n * (18 + 24) / 1e9;
    printf("bits %2d threads %4d items %2d tile %5d mode %d occ %d smem %6zu | hist+scan %.3f ms  scatter %.3f ms = %.0f GB/s (42 B/event)  per 1e9 events: %.2f ms  misplaced %u\n",
           BITS, THREADS, ITEMS, TILE, MODE, occ, smem, best_h, best_s, gb / (best_s * 1e-3), best_s * 1e9 / (double)n, bad);
    fflush(stdout);
}

int main(int argc, char** argv) {
    const u64 n = argc > 1 ? (u64)atof(argv[1]) : (1ull << 28);
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    Bufs B;
    const u64 c = n + 8192;
    CK(cudaMalloc(&B.hi, c * 8)); CK(cudaMalloc(&B.lo, c * 8)); CK(cudaMalloc(&B.ph, c)); CK(cudaMalloc(&B.ty, c));
    CK(cudaMalloc(&B.out, c * 24));
    CK(cudaMalloc(&B.hist, (u64)sms * 8 * 2048 * 4)); CK(cudaMalloc(&B.off, (u64)sms * 8 * 2048 * 4));
    CK(cudaMalloc(&B.bin_start, 2048 * 4)); CK(cudaMalloc(&B.gcursor, 2048 * 4)); CK(cudaMalloc(&B.ticket, 4)); CK(cudaMalloc(&B.bad, 4));
    gen_kernel<<<sms * 8, 256>>>(B.hi, B.lo, B.ph, B.ty, n, n / 100);
    CK(cudaDeviceSynchronize());
    printf("events %llu, pods %llu, SMs %d\n", n, n / 100, sms);
#define BOTH(BITS, T, I) run<BITS, T, I, 0>(B, n, sms); run<BITS, T, I, 1>(B, n, sms);
    BOTH(8, 256, 8) BOTH(8, 512, 8) BOTH(8, 384, 16)
    BOTH(9, 256, 8) BOTH(9, 512, 8)
    BOTH(10, 256, 8) BOTH(10, 512, 8) BOTH(10, 512, 12) BOTH(10, 256, 16)
    BOTH(11, 256, 8) BOTH(11, 512, 8) BOTH(11, 512, 12) BOTH(11, 1024, 8)
    return 0;
}
