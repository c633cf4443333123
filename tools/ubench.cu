// Micro-benchmarks of the warp-level primitives a radix ranking step can be built from (B200).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench tools/ubench.cu ; run on the GPU box.
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned int u32;
#define ITER 2048
template <int MODE>
__global__ void __launch_bounds__(256) k(u32* out, u32 seed) {
    __shared__ u32 sh[8][256];
    u32 x = threadIdx.x * 2654435761u + blockIdx.x * 40503u + seed;
    u32 acc = 0;
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 8 * 256; i += 256) (&sh[0][0])[i] = 0;
    __syncthreads();
#pragma unroll 4
    for (int it = 0; it < ITER; it++) {
        x = x * 1664525u + 1013904223u;
        u32 d = (x >> 13) & 255;
        if (MODE == 0) acc += __match_any_sync(0xffffffffu, d);
        if (MODE == 1) {
            u32 peers = 0xffffffffu;
#pragma unroll
            for (int b = 0; b < 8; b++) { bool bit = (d >> b) & 1; u32 m = __ballot_sync(0xffffffffu, bit); peers &= bit ? m : ~m; }
            acc += peers;
        }
        if (MODE == 2) acc += __ballot_sync(0xffffffffu, d & 1);
        if (MODE == 3) acc += __shfl_sync(0xffffffffu, d, (lane + 1) & 31);
        if (MODE == 4) acc += atomicAdd(&sh[warp][d], 1u);
        if (MODE == 5) acc += __reduce_add_sync(0xffffffffu, d);
        if (MODE == 6) { atomicAdd(&sh[warp][d], 1u); }
        if (MODE == 7) acc += __popc(d * x) + (d ^ x);
        if (MODE == 8) {   // 4-bit match: 4 ballots
            u32 peers = 0xffffffffu;
#pragma unroll
            for (int b = 0; b < 4; b++) { bool bit = (d >> b) & 1; u32 m = __ballot_sync(0xffffffffu, bit); peers &= bit ? m : ~m; }
            acc += peers;
        }
        if (MODE == 9) { sh[warp][lane] = d; __syncwarp(); acc += sh[warp][(lane + 5) & 31]; __syncwarp(); }
    }
    out[blockIdx.x * 256 + threadIdx.x] = acc + sh[warp][lane];
}
template <int MODE>
void run(const char* name, u32* d_out) {
    int blocks = 148 * 8;
    k<MODE><<<blocks, 256>>>(d_out, 1);
    cudaDeviceSynchronize();
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a);
    k<MODE><<<blocks, 256>>>(d_out, 2);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    double warp_ops = (double)blocks * 8 * ITER;
    double cyc = ms * 1e-3 * 1.965e9;   // at max clock
    printf("%-28s %8.3f ms  %6.2f cycles per warp-op per SM\n", name, ms, cyc / (warp_ops / 148));
}
int main() {
    u32* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    run<7>("alu baseline (loop body)", d);
    run<0>("match.any 8-bit", d);
    run<1>("8 ballots + lop", d);
    run<8>("4 ballots + lop", d);
    run<2>("1 ballot", d);
    run<3>("shfl", d);
    run<5>("redux.add", d);
    run<4>("smem atomicAdd (ret)", d);
    run<6>("smem red (no ret)", d);
    run<9>("sts+syncwarp+lds+syncwarp", d);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
