// pte_partition_u64_static: one radix scatter pass for callers that are not allowed to allocate.
//
// Who that is: code that runs inside somebody else's memory budget.  An inference server or a
// database engine that hands every byte of device memory to its own pool at start-up leaves
// nothing for cudaMalloc to give out later, and treats a library that allocates behind its back
// as a bug; a plug-in called from a latency-critical loop cannot afford the implicit device
// synchronisation that cudaMalloc/cudaFree bring.  Such a caller already owns the input and the
// output buffer.  What it cannot supply is the sort's scratch -- digit histogram, tile ticket and
// look-back words -- so that lives here, in a fixed array in the library's device image.
//
// The price: the array has a fixed size (so the call has a size limit), every GPU that loads the
// library carries it until the process exits, and all calls on one GPU share it.
#include "../../include/pte.h"
#include "pte_radix.cuh"

#include <cstdio>
#include <cstring>
#include <mutex>

namespace {

constexpr int kThreads = 512, kItems = 8, kTileRecords = kThreads * kItems;
constexpr u64 kMaxRecords = PTE_STATIC_MAX_RECORDS;
constexpr u32 kMaxTiles = (u32)((kMaxRecords + kTileRecords - 1) / kTileRecords);

// words: [0,256) digit histogram | [256] tile ticket | [512, 512 + tiles*256) look-back, one word per (tile, digit)
constexpr u32 kHistOff = 0, kTicketOff = 256, kLookbackOff = 512;
constexpr u32 kScratchWords = kLookbackOff + kMaxTiles * kRadix;

// The scratch itself (about 1 MB).  In host code this name is only a handle for the runtime,
// never an address a kernel could use: the address comes from cudaGetSymbolAddress below.
static __device__ __align__(256) u32 g_static_scratch[kScratchWords];

struct DeviceState {
    bool ready = false;
    u32* scratch = nullptr;   // device address of g_static_scratch on this GPU
    int sms = 0, occ = 0;
};
constexpr int kMaxDevices = 64;
DeviceState g_state[kMaxDevices];
std::mutex g_mu;

// digit counts of one pass; scalar loads only, so the caller's buffer needs no padding
__global__ void __launch_bounds__(256)
static_hist_kernel(const u64* __restrict__ in, u64 n, int shift, u32 mask, u32* __restrict__ hist) {
    __shared__ u32 sh[kRadix];
    sh[threadIdx.x] = 0;
    __syncthreads();
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x)
        atomicAdd(&sh[(u32)(in[i] >> shift) & mask], 1u);
    __syncthreads();
    if (sh[threadIdx.x]) atomicAdd(&hist[threadIdx.x], sh[threadIdx.x]);
}

int fail(int code, char* err, size_t cap, const char* msg) {
    if (err && cap) snprintf(err, cap, "%s", msg);
    return code;
}

}  // namespace

extern "C" int pte_partition_u64_static(const uint64_t* in, uint64_t* out, uint64_t n, int bit_lo, int n_bits,
                                        void* stream, char* err, size_t err_cap) {
    char msg[256];
    if (err && err_cap) err[0] = 0;
    if (n_bits < 1 || n_bits > 8 || bit_lo < 0 || bit_lo + n_bits > 64)
        return fail(PTE_ERR_INVALID, err, err_cap, "the digit must be 1..8 bits wide and lie inside the 64-bit record");
    if (n > kMaxRecords) {
        snprintf(msg, sizeof msg, "%llu records exceed PTE_STATIC_MAX_RECORDS (%llu): the built-in scratch array "
                 "cannot cover more tiles; split the input or use an engine handle",
                 (unsigned long long)n, (unsigned long long)kMaxRecords);
        return fail(PTE_ERR_CAPACITY, err, err_cap, msg);
    }
    if (n == 0) return PTE_OK;
    if (!in || !out || in == out) return fail(PTE_ERR_INVALID, err, err_cap, "in and out must be two distinct buffers");

    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
        cudaGetLastError();
        return fail(PTE_ERR_NO_DEVICE, err, err_cap, "no current CUDA device");
    }
    DeviceState st;
    {
        std::lock_guard<std::mutex> lock(g_mu);
        DeviceState& d = g_state[dev];
        if (!d.ready) {
            // once per GPU, with that GPU current: turn the symbol into an address kernels can take
            e = cudaGetSymbolAddress((void**)&d.scratch, g_static_scratch);
            if (e == cudaSuccess)
                e = cudaFuncSetAttribute(radix_pass_kernel<kThreads, kItems>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)sizeof(RadixPassSmem<kThreads, kItems>));
            if (e == cudaSuccess)
                e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.occ, radix_pass_kernel<kThreads, kItems>, kThreads,
                                                                  sizeof(RadixPassSmem<kThreads, kItems>));
            if (e == cudaSuccess) e = cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, dev);
            if (e != cudaSuccess || d.occ < 1 || !d.scratch) {
                snprintf(msg, sizeof msg, "set-up of the static scratch failed: %s", cudaGetErrorString(e));
                cudaGetLastError();
                return fail(PTE_ERR_CUDA, err, err_cap, msg);
            }
            d.ready = true;
        }
        st = d;
    }

    cudaStream_t s = (cudaStream_t)stream;
    const u64* src = reinterpret_cast<const u64*>(in);
    u64* dst = reinterpret_cast<u64*>(out);
    const u32 tiles = (u32)((n + kTileRecords - 1) / kTileRecords);
    const u32 mask = (1u << n_bits) - 1;
    // the array still holds the previous call's counts and flags: clear the part this call reads
    e = cudaMemsetAsync(st.scratch, 0, sizeof(u32) * (kLookbackOff + (size_t)tiles * kRadix), s);
    if (e == cudaSuccess) {
        int hgrid = (int)((n + 2047) / 2048);
        if (hgrid > st.sms * 8) hgrid = st.sms * 8;
        static_hist_kernel<<<hgrid, 256, 0, s>>>(src, n, bit_lo, mask, st.scratch + kHistOff);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) {
        radix_pass_kernel<kThreads, kItems><<<st.sms * st.occ, kThreads, sizeof(RadixPassSmem<kThreads, kItems>), s>>>(
            src, dst, nullptr, n, bit_lo, mask, st.scratch + kHistOff, st.scratch + kLookbackOff, st.scratch + kTicketOff);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) {
        snprintf(msg, sizeof msg, "static partition failed: %s", cudaGetErrorString(e));
        return fail(PTE_ERR_CUDA, err, err_cap, msg);
    }
    return PTE_OK;
}
