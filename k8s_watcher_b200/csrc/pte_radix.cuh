// LSD radix sort of 64-bit records on an arbitrary bit field, 8 bits per pass.
//
//   radix_hist_kernel  one read of the records -> digit counts of every pass
//   radix_pass_kernel  one scatter pass: read 8 B, write 8 B per record, single sweep.
//                      Tiles are chained with a decoupled look-back over per-tile digit
//                      counts, digits are ranked inside a warp with peer masks built by shared-memory atomicOr + popc
//                      (stable: lower lanes first), and every tile is re-ordered in shared
//                      memory first so that global writes are runs of consecutive records.
//
// No reference counterpart: the reference handles one event at a time in Python
// (watcher/pod_watcher.py:266-269); sorting by (uid key, rv) is how the bulk engine
// brings each pod's events together (bulk-equivalent predicate).
#pragma once
#include "pte_common.cuh"

static constexpr int kRadix = 256;
static constexpr int kMaxPasses = 8;
static constexpr u32 kLbPartial = 1u << 30, kLbInclusive = 2u << 30, kLbValue = (1u << 30) - 1;

__device__ __forceinline__ u32 ld_relaxed_u32(const u32* p) {
    u32 v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u32(u32* p, u32 v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ u32 lanemask_lt() {
    u32 m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// Exclusive scan of one value per thread over the first 256 threads of the block.
// Must be called by exactly threads 0..255 (named barrier 1), `tot` = 8 words of smem.
__device__ __forceinline__ u32 excl_scan_256(u32 v, u32* tot, u32* total_out) {
    const u32 lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    u32 inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        u32 t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) tot[w] = inc;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    u32 base = 0, all = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        u32 t = tot[k];
        if (k < (int)w) base += t;
        all += t;
    }
    if (total_out) *total_out = all;
    asm volatile("bar.sync 1, 256;" ::: "memory");   // tot may be reused by the caller
    return base + inc - v;
}

// ---- digit counts for n_passes passes starting at bit shift0 (field ends at bit_hi) ----
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
radix_hist_kernel(const u64* __restrict__ in, const u64* __restrict__ n_ptr, u64 n_fixed, int shift0,
                  int n_passes, int bit_hi, u32* __restrict__ hist) {
    __shared__ u32 sh[kMaxPasses * kRadix];
    for (int k = threadIdx.x; k < n_passes * kRadix; k += THREADS) sh[k] = 0;
    __syncthreads();
    const u64 n = n_ptr ? *n_ptr : n_fixed;
    u32 masks[kMaxPasses];
#pragma unroll
    for (int p = 0; p < kMaxPasses; p++) {
        int rem = bit_hi - (shift0 + 8 * p);
        masks[p] = rem >= 8 ? 0xFFu : (rem > 0 ? (1u << rem) - 1 : 0u);
    }
    const u64 stride = (u64)gridDim.x * THREADS * 2;
    for (u64 i = ((u64)blockIdx.x * THREADS + threadIdx.x) * 2; i < n; i += stride) {
        ulonglong2 v = *reinterpret_cast<const ulonglong2*>(in + i);   // buffers are padded: i+1 is readable
        const bool two = i + 1 < n;
#pragma unroll
        for (int p = 0; p < kMaxPasses; p++) {
            if (p < n_passes) {
                int s = shift0 + 8 * p;
                atomicAdd(&sh[p * kRadix + ((u32)(v.x >> s) & masks[p])], 1u);
                if (two) atomicAdd(&sh[p * kRadix + ((u32)(v.y >> s) & masks[p])], 1u);
            }
        }
    }
    __syncthreads();
    for (int k = threadIdx.x; k < n_passes * kRadix; k += THREADS)
        if (sh[k]) atomicAdd(&hist[k], sh[k]);
}

// ---- one scatter pass ----
template <int THREADS, int ITEMS>
struct RadixPassSmem {
    static constexpr int WARPS = THREADS / 32;
    static constexpr int TILE = THREADS * ITEMS;
    u64 stage[TILE];
    u32 warp_hist[WARPS][kRadix];   // per warp: digit counts, then running offsets inside the tile
    u32 peer_mask[WARPS][kRadix];   // per warp and digit: lanes of the step in flight holding that digit (0 between steps)
    u32 bin_base[kRadix];
    u32 dst_off[kRadix];
    u32 tot[8];
    u32 tile, next_tile;
};

// One digit of one warp-wide step.  Lanes holding the same digit have to find each other (stable
// ranking: lower lanes first); the first of them then advances the warp's running offset for the
// digit.  Three ways to get the peer mask on sm_100a (tools/ubench.cu, profiles/):
//   MATCH.ANY        one instruction, but 62 cycles per warp instruction per SM on the ADU pipe;
//   eight ballots    ~22 cycles of the vote path plus ~40 ALU-pipe instructions;
//   shared atomicOr  every lane ORs its lane bit into the warp's mask word of its digit and reads the
//                    word back: under 3 cycles per op, but ~3.4 bank-conflict wavefronts each.
// The pass uses the atomicOr form.  BALLOT_ITEMS > 0 ranks that many of a thread's records with ballots
// instead (the two forms load different pipes), kept as a tuning option: every mix measured slower than
// all-atomic on B200 (profiles/README.md).
template <bool CHECK_VALID>
__device__ __forceinline__ u32 rank_by_ballot(u32* __restrict__ warp_off, u32 d, bool valid, u32 lt) {
    u32 peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < 8; b++) {
        const bool bit = (d >> b) & 1u;
        const u32 m = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? m : ~m;
    }
    if (CHECK_VALID) {
        const u32 vm = __ballot_sync(0xffffffffu, valid);
        peers &= valid ? vm : ~vm;
    }
    const u32 before = peers & lt;
    u32 base = 0;
    if (before == 0 && (!CHECK_VALID || valid)) base = atomicAdd(warp_off + d, (u32)__popc(peers));
    base = __shfl_sync(0xffffffffu, base, (__ffs(peers) - 1) & 31);
    return base + __popc(before);
}

template <bool CHECK_VALID>
__device__ __forceinline__ u32 rank_by_atomic(u32* __restrict__ warp_off, u32* __restrict__ peer_mask, u32 d, bool valid,
                                              u32 lt, u32 lane_bit) {
    const bool on = !CHECK_VALID || valid;
    if (on) atomicOr(peer_mask + d, lane_bit);
    __syncwarp();
    const u32 peers = on ? peer_mask[d] : 0u;     // __syncwarp orders it after every peer's atomicOr
    __syncwarp();
    const u32 before = peers & lt;
    u32 base = 0;
    if (before == 0 && on) {
        base = atomicAdd(warp_off + d, (u32)__popc(peers));
        peer_mask[d] = 0;
    }
    __syncwarp();
    base = __shfl_sync(0xffffffffu, base, (__ffs(peers) - 1) & 31);
    return base + __popc(before);
}

template <int THREADS, int ITEMS, int BALLOT_ITEMS = 0>
__global__ void __launch_bounds__(THREADS, (THREADS <= 256 ? 4 : (THREADS <= 512 ? 2 : 1)))
radix_pass_kernel(const u64* __restrict__ in, u64* __restrict__ out, const u64* __restrict__ n_ptr, u64 n_fixed,
                  int shift, u32 mask, const u32* __restrict__ hist, u32* __restrict__ lookback,
                  u32* __restrict__ ticket) {
    static_assert(THREADS >= 256 && THREADS % 32 == 0, "needs one thread per digit");
    using S = RadixPassSmem<THREADS, ITEMS>;
    constexpr int WARPS = S::WARPS, TILE = S::TILE;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    S& s = *reinterpret_cast<S*>(smem_raw);

    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const u32 lt = lanemask_lt();
    const u32 lane_bit = 1u << lane;
    const u64 n = n_ptr ? *n_ptr : n_fixed;
    const u32 wbase = warp * 32 * ITEMS + lane;

    if (tid < kRadix) s.bin_base[tid] = excl_scan_256(hist[tid], s.tot, nullptr);
    for (int k = tid; k < WARPS * kRadix; k += THREADS) (&s.peer_mask[0][0])[k] = 0;
    if (tid == 0) s.tile = atomicAdd(ticket, 1u);
    __syncthreads();

    // warp-striped load: warp w owns records [w*32*ITEMS, (w+1)*32*ITEMS) of its tile.  The registers are
    // free again as soon as a tile's records sit in shared memory, so the NEXT tile is loaded right then
    // and its DRAM latency hides behind this tile's look-back and write-out.
    u64 key[ITEMS];
    u32 tile = s.tile;
    auto load_tile = [&](u32 t) {
        const u64 base = (u64)t * TILE;
#pragma unroll
        for (int k = 0; k < ITEMS; k++) {
            const u64 i = base + wbase + k * 32;
            key[k] = i < n ? __ldcs(in + i) : ~0ull;
        }
    };
    load_tile(tile);

    for (;;) {
        const u64 tile_base = (u64)tile * TILE;
        if (tile_base >= n) break;
        const u32 tile_valid = (u32)min((u64)TILE, n - tile_base);
        const bool full_tile = tile_valid == TILE;
        for (int k = tid; k < WARPS * kRadix; k += THREADS) (&s.warp_hist[0][0])[k] = 0;
        __syncthreads();   // counters cleared; previous tile fully written out of `stage`

        // (2) early counts: a cheap pass of shared-memory reductions, so that the tile's digit counts
        // reach the look-back chain long before the ranking is done
#pragma unroll
        for (int k = 0; k < ITEMS; k++)
            if (full_tile || wbase + k * 32 < tile_valid) atomicAdd(&s.warp_hist[warp][(u32)(key[k] >> shift) & mask], 1u);
        if (tid == 0) s.next_tile = atomicAdd(ticket, 1u);
        __syncthreads();

        // (3) per digit: tile count (published at once), each warp's start inside the tile
        u32 count = 0, tile_excl = 0;
        if (tid < kRadix) {
            u32 run = 0;
#pragma unroll
            for (int w = 0; w < WARPS; w++) run += s.warp_hist[w][tid];
            count = run;
            st_relaxed_u32(lookback + (u64)tile * kRadix + tid, (tile == 0 ? kLbInclusive : kLbPartial) | count);
            tile_excl = excl_scan_256(count, s.tot, nullptr);
            run = tile_excl;
#pragma unroll
            for (int w = 0; w < WARPS; w++) {
                u32 t = s.warp_hist[w][tid];
                s.warp_hist[w][tid] = run;
                run += t;
            }
        }
        __syncthreads();
        const u32 next_tile = s.next_tile;

        // (4) rank inside the warp, in record order, and drop every record at its place in tile order
        if (full_tile) {
#pragma unroll
            for (int k = 0; k < ITEMS; k++) {
                const u32 d = (u32)(key[k] >> shift) & mask;
                const u32 pos = k < BALLOT_ITEMS ? rank_by_ballot<false>(s.warp_hist[warp], d, true, lt)
                                                 : rank_by_atomic<false>(s.warp_hist[warp], s.peer_mask[warp], d, true, lt, lane_bit);
                s.stage[pos] = key[k];
            }
        } else {
#pragma unroll
            for (int k = 0; k < ITEMS; k++) {
                const bool valid = wbase + k * 32 < tile_valid;
                const u32 d = (u32)(key[k] >> shift) & mask;
                const u32 pos = k < BALLOT_ITEMS ? rank_by_ballot<true>(s.warp_hist[warp], d, valid, lt)
                                                 : rank_by_atomic<true>(s.warp_hist[warp], s.peer_mask[warp], d, valid, lt, lane_bit);
                if (valid) s.stage[pos] = key[k];
            }
        }
        load_tile(next_tile);      // in flight during (5) and (6)

        // (5) decoupled look-back: records of this digit in all earlier tiles.  Four predecessors are
        // read per step so that the walk costs one L2 round trip per four tiles.
        if (tid < kRadix) {
            u32 excl = 0;
            if (tile > 0) {
                int t = (int)tile - 1;
                bool done = false;
                while (!done) {
                    u32 v[4];
#pragma unroll
                    for (int j = 0; j < 4; j++)
                        v[j] = t - j >= 0 ? ld_relaxed_u32(lookback + (u64)(t - j) * kRadix + tid) : kLbInclusive;
                    int used = 0;
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        if (done || (v[j] >> 30) == 0) break;
                        excl += v[j] & kLbValue;
                        used++;
                        if ((v[j] >> 30) == 2) done = true;
                    }
                    t -= used;
                    if (!done && used == 0) __nanosleep(40);
                }
                st_relaxed_u32(lookback + (u64)tile * kRadix + tid, kLbInclusive | (excl + count));
            }
            s.dst_off[tid] = s.bin_base[tid] + excl - tile_excl;
        }
        __syncthreads();

        // (6) coalesced runs out to global memory
#pragma unroll
        for (int k = 0; k < ITEMS; k++) {
            u32 i = tid + k * THREADS;
            if (i < tile_valid) {
                u64 v = s.stage[i];
                out[s.dst_off[(u32)(v >> shift) & mask] + i] = v;
            }
        }
        tile = next_tile;
    }
}
