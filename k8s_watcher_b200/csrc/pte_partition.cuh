// Partition-first replay: the T-trace of a large log without a global dictionary walk and without a sort.
//
// The per-event rule (it would sit at watcher/pod_watcher.py:236) only ever looks at
// the PREVIOUS event of the same uid (metadata.uid, pod_watcher.py:164).  So the log is first split by a
// hash of the uid into 1024 partitions, each keeping arrival order (one stable scatter of 24-byte records).
// A partition holds 1/1024 of the pods: its slice of the uid dictionary is a few hundred KB and stays in
// L2 while it is probed, and the last-known phase of all its pods fits in the shared memory of one CTA.
// The CTA then simply replays the partition in order.
//
//   part_hist_kernel     uid hash -> partition counts (+ replay-order check)          16 B read / event
//   part_scan_kernel     partition starts, segment table (long partitions are cut into segments)
//   part_scatter_kernel  events -> 24-byte records (uid_hi, uid_lo, pos|type|phase), grouped by partition,
//                        arrival order kept: warp-striped tile, ballot ranking, per-warp counts, tile
//                        re-ordered in shared memory, runs written out; tiles chained by decoupled look-back
//                                                                                     18 B read + 24 B written
//   part_replay_kernel   one CTA per segment: tiles come in by 1-D bulk async copy (cp.async.bulk + mbarrier,
//                        double-buffered), uids are interned in the partition's dictionary slice (L2 hits),
//                        events are dealt to the warp that owns their slot (slot mod 16) keeping order, each
//                        warp replays its queue 32 events at a time against the state bytes in shared
//                        memory, emitted positions flip bits of the emit bitmap                24 B read
//   pending_fixup_kernel first events of a uid inside a later segment of a cut partition: their
//                        predecessor state is looked up in the state tables the earlier segments exported
//
// Skew (a pod that owns a few per cent of all events makes its partition many times the mean) is what the
// segments are for: every segment replays independently from an "unknown" state, and only the first event
// of each uid in a non-first segment needs the fix-up.
#pragma once
#include "pte_kernels.cuh"

// Measured on B200, 1 B events (profiles/r02_partition): ballots + word-by-word write-out 18.2 ms; MATCH.ANY ranking 20.9 ms
// (the ADU pipe serialises it); one lane per record in the write-out 22.5 ms (three times the L2 write transactions); both 24.9 ms.
#ifndef PTE_SCATTER_WORDWISE
#define PTE_SCATTER_WORDWISE 1
#endif
#ifndef PTE_SCATTER_MATCH
#define PTE_SCATTER_MATCH 0
#endif
static constexpr int kPartBits = 10;
static constexpr int kParts = 1 << kPartBits;
static constexpr u64 kPartSalt = 0x7061727469746E21ull;
static constexpr u32 kMinSegment = 1u << 15;          // records; shorter partitions are never cut
static constexpr u32 kMaxSegments = 16 * kParts;
static constexpr u32 kSampleSegments = 4;             // extra entries behind the segment table: the polarity sample

// partition of a uid: one xor-shift and one multiply of uid_lo (the counting pass reads that column only)
__host__ __device__ __forceinline__ u32 part_of(u64 uid_lo) {
    u64 x = uid_lo ^ kPartSalt;
    x ^= x >> 32;
    return (u32)((x * 0x9E3779B97F4A7C15ull) >> (64 - kPartBits));
}
// 24 more hash bits, of both halves, computed once by the scatter and carried in bits [38, 62) of the record's third word:
// the replay takes the uid's home line and its 8-bit fingerprint from them
__host__ __device__ __forceinline__ u32 uid_hash24(u64 hi, u64 lo) {
    u64 x = lo ^ (hi * 0xD1B54A32D192ED03ull);
    x ^= x >> 29;
    return (u32)((x * 0xBF58476D1CE4E5B9ull) >> 40);
}

struct __align__(16) Segment {
    u32 beg, end;          // records [beg, end) of the partition-ordered buffer
    u32 part;              // partition
    u32 ordinal;           // 0 for the first segment of its partition; bit 31: another segment follows
};
struct __align__(16) PendingHead {
    u64 meta;              // pos << 8 | type << 6 | effective phase
    u32 seg;               // global segment index (its partition's earlier segments are seg-1, seg-2, ...)
    u32 slot;              // slot inside the partition's dictionary slice
};

enum : u32 { kFlagPendingFull = 8u, kFlagSegmentsFull = 16u };

// state byte of a pod: 0 = no live phase known (never seen, or last event DELETED), 1 + phase, 0xFF = not seen in this segment
__device__ __forceinline__ u32 state_after(u32 tp) { return (tp & 0x80u) ? 0u : 1u + (tp & 0x3Fu); }
__device__ __forceinline__ bool emits(u32 pred_state, u32 tp) {
    return (tp & 0x80u) || pred_state == 0u || pred_state - 1u != (tp & 0x3Fu);
}

// exclusive scan over all threads of a block of up to 1024 threads; `part` = 32 words of shared memory
__device__ __forceinline__ u32 block_excl_scan(u32 v, u32* part, u32* total) {
    const u32 lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    u32 inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const u32 t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= (u32)o) inc += t; }
    if (lane == 31) part[w] = inc;
    __syncthreads();
    u32 base = 0, all = 0;
    for (u32 k = 0; k < nw; k++) { const u32 t = part[k]; if (k < w) base += t; all += t; }
    if (total) *total = all;
    __syncthreads();
    return base + inc - v;
}

// ---- partition counts ---------------------------------------------------------------------------------
struct PartHistArgs {
    const u64* uid_lo; const u64* rv; const u32* keep;     // keep: FILTERED only (filter_flag_kernel's bitmap)
    u64 n; u32* hist; Counters* ctr;
};

template <bool FILTERED>
__global__ void __launch_bounds__(256) part_hist_kernel(PartHistArgs a) {
    __shared__ u32 sh[kParts];
    for (int k = threadIdx.x; k < kParts; k += 256) sh[k] = 0;
    __syncthreads();
    const u32 lane = threadIdx.x & 31;
    u32 flags = 0;
    const u64 n4 = (a.n + 3) / 4;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (u64 wb = (u64)blockIdx.x * blockDim.x + (threadIdx.x & ~31u); wb < n4; wb += stride) {
        const u64 i0 = (wb + lane) * 4;
        const ulonglong2* p = reinterpret_cast<const ulonglong2*>(a.uid_lo + i0);     // padded: readable past n
        const ulonglong2 lx = __ldcs(p), ly = __ldcs(p + 1);
        u32 keep_bits = 0xFu;
        if (FILTERED) keep_bits = (a.keep[i0 >> 5] >> (i0 & 31)) & 0xFu;
        else {
            // replay order check: rv must not decrease along arrival order (filter_flag_kernel does it for FILTERED)
            const ulonglong2* q = reinterpret_cast<const ulonglong2*>(a.rv + i0);
            const ulonglong2 x = __ldcs(q), y = __ldcs(q + 1);
            u64 prev = __shfl_up_sync(0xffffffffu, y.y, 1);
            if (lane == 0) prev = i0 > 0 ? a.rv[i0 - 1] : 0;
            if ((i0 < a.n && x.x < prev) || (i0 + 1 < a.n && x.y < x.x) || (i0 + 2 < a.n && y.x < x.y) ||
                (i0 + 3 < a.n && y.y < y.x))
                flags |= kFlagUnordered;
        }
        const u64 lo[4] = {lx.x, lx.y, ly.x, ly.y};
#pragma unroll
        for (int k = 0; k < 4; k++)
            if (i0 + k < a.n && ((keep_bits >> k) & 1u)) atomicAdd(&sh[part_of(lo[k])], 1u);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < kParts; k += 256)
        if (sh[k]) atomicAdd(&a.hist[k], sh[k]);
    flags = __reduce_or_sync(0xffffffffu, flags);
    if (lane == 0 && flags) atomicOr(&a.ctr->flags, flags);
}

// ---- partition starts + segment table (one block of kParts threads) --------------------------------------
// Segment length: seg_q8 / 256 of the mean partition size, never below kMinSegment.
__global__ void __launch_bounds__(kParts)
part_scan_kernel(const u32* __restrict__ hist, u32* __restrict__ part_start, Segment* __restrict__ segs, u32* __restrict__ n_segs,
                 Counters* __restrict__ ctr, u32 seg_q8, int set_survived) {
    __shared__ u32 s_part[32];
    const u32 t = threadIdx.x;
    const u32 c = hist[t];
    u32 total;
    const u32 ex = block_excl_scan(c, s_part, &total);
    part_start[t] = ex;
    if (t == 0) { part_start[kParts] = total; if (set_survived) ctr->n_survived = total; }
    const u64 want = ((u64)total * seg_q8 + (u64)kParts * 256 - 1) / ((u64)kParts * 256);
    const u32 seg_len = (u32)max((u64)kMinSegment, want);
    const u32 ns = c ? (c + seg_len - 1) / seg_len : 0;
    u32 stotal;
    const u32 sb = block_excl_scan(ns, s_part, &stotal);
    if (t == 0) {
        n_segs[1] = kSampleSegments;
        *n_segs = min(stotal, kMaxSegments);
        if (stotal > kMaxSegments) atomicOr(&ctr->flags, kFlagSegmentsFull);
    }
    for (u32 j = 0; j < ns && sb + j < kMaxSegments; j++) {
        Segment s;
        s.beg = ex + j * seg_len;
        s.end = min(ex + c, s.beg + seg_len);
        s.part = t;
        s.ordinal = j | (j + 1 < ns ? 0x80000000u : 0u);
        segs[sb + j] = s;
    }
    // the polarity sample: a prefix of the first partitions, replayed without side effects on the bitmap
    if (t < kSampleSegments) {
        Segment s;
        s.beg = ex; s.end = ex + min(c, 1u << 16); s.part = t; s.ordinal = 0;
        segs[kMaxSegments + t] = s;
    }
}

// ---- stable scatter of the events into partition order -------------------------------------------------------
template <int THREADS, int ITEMS>
struct ScatterCfg {
    static constexpr int kWarps = THREADS / 32, kTile = THREADS * ITEMS, kWarpRecs = 32 * ITEMS;
    static constexpr int kWords = kParts / 2;                 // u16 pairs: one word covers two partitions
    static constexpr int kWordsPerThread = kWords / THREADS;  // words a thread owns in everything that is per partition
    static_assert(kWords % THREADS == 0 && kWordsPerThread >= 1, "threads must divide kParts / 2");
    static_assert(kTile % 16 == 0 && kTile / 16 <= THREADS, "phase/type bytes come in as one 16-byte load per thread");
    static_assert(kWarpRecs * 1 < 65536 && kTile < 65536, "u16 counters");
};

template <int THREADS, int ITEMS>
struct ScatterSmem {
    using C = ScatterCfg<THREADS, ITEMS>;
    u64 stage[C::kTile * 3];                 // the tile in partition order, 3 words per record
    u32 whist[C::kWarps][C::kWords];         // per warp and partition: count, then start inside the tile's run (u16 pairs)
    u32 tile_excl[kParts];                   // start of each partition's run inside the tile
    u32 dst_off[kParts];                     // global record index of the run minus tile_excl
    unsigned short sbin[C::kTile];           // partition of every staged record
    u8 ph[C::kTile], ty[C::kTile];
    u32 scan_part[32];
    u32 tile, next_tile, survivors;
};

struct ScatterArgs {
    const u64* uid_hi; const u64* uid_lo; const u8* phase; const u8* type; const u32* keep;
    u64 n;
    const u32* part_start;
    u64* out;                                // 3 x u64 per survivor
    u32* lookback; u32* ticket;
    Counters* ctr;
};

template <int THREADS, int ITEMS, int MIN_BLOCKS, bool FILTERED>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS) part_scatter_kernel(ScatterArgs a) {
    using C = ScatterCfg<THREADS, ITEMS>;
    constexpr int kTile = C::kTile, kWarps = C::kWarps, kWarpRecs = C::kWarpRecs, WPT = C::kWordsPerThread;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    ScatterSmem<THREADS, ITEMS>& s = *reinterpret_cast<ScatterSmem<THREADS, ITEMS>*>(smem_raw);
    unsigned short* whist16 = reinterpret_cast<unsigned short*>(&s.whist[0][0]);
    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const u32 lt = lanemask_lt();
    const u64 n = a.n;
    // thread t owns the partitions of words t, t + THREADS, ... (two partitions per word)
    u32 pbase[WPT][2];
#pragma unroll
    for (int j = 0; j < WPT; j++) {
        const u32 w = tid + j * THREADS;
        pbase[j][0] = a.part_start[2 * w]; pbase[j][1] = a.part_start[2 * w + 1];
    }
    u32 del_surv = 0, flags = 0;

    if (tid == 0) s.tile = atomicAdd(a.ticket, 1u);
    __syncthreads();
    u32 tile = s.tile;

    u64 hi[ITEMS], lo[ITEMS];
    uint4 phv = make_uint4(0, 0, 0, 0), tyv = make_uint4(0, 0, 0, 0);
    u32 keepw[FILTERED ? ITEMS : 1];
    auto load_tile = [&](u32 t) {
        const u64 base = (u64)t * kTile;
        if (base >= n) return;
#pragma unroll
        for (int k = 0; k < ITEMS; k++) {
            const u64 i = base + warp * kWarpRecs + k * 32 + lane;       // buffers are padded by more than a tile
            lo[k] = __ldcs(a.uid_lo + i);
            hi[k] = __ldcs(a.uid_hi + i);
            if (FILTERED) keepw[k] = __ldg(a.keep + (i >> 5));             // one word per warp step
        }
        if (tid < kTile / 16) {
            phv = __ldcs(reinterpret_cast<const uint4*>(a.phase + base) + tid);
            tyv = __ldcs(reinterpret_cast<const uint4*>(a.type + base) + tid);
        }
    };
    load_tile(tile);

    for (;;) {
        const u64 tile_base = (u64)tile * kTile;
        if (tile_base >= n) break;
        const u32 tile_valid = (u32)min((u64)kTile, n - tile_base);
        // (1) phase/type bytes of the tile, per-warp counters cleared
        if (tid < kTile / 16) {
            reinterpret_cast<uint4*>(s.ph)[tid] = phv;
            reinterpret_cast<uint4*>(s.ty)[tid] = tyv;
        }
        for (int k = tid; k < kWarps * C::kWords; k += THREADS) (&s.whist[0][0])[k] = 0;
        if (tid == 0) s.next_tile = atomicAdd(a.ticket, 1u);
        __syncthreads();

        // (2) rank inside the warp, in arrival order: ballots over the partition bits give the lanes of the
        // step that go to the same partition; the first of them advances the warp's counter for it
        u32 bin[ITEMS];
        unsigned short lrank[ITEMS];
#pragma unroll
        for (int k = 0; k < ITEMS; k++) {
            const u32 j = warp * kWarpRecs + k * 32 + lane;
            bool valid = j < tile_valid;
            if (FILTERED) valid = valid && ((keepw[k] >> lane) & 1u);
            const u32 b = part_of(lo[k]);
            // lanes of this step that go to the same partition
#if PTE_SCATTER_MATCH
            u32 peers = __match_any_sync(0xffffffffu, valid ? b : 0xFFFF0000u | lane);
#else
            u32 peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
            for (int q = 0; q < kPartBits; q++) {
                const bool bit = (b >> q) & 1u;
                const u32 m = __ballot_sync(0xffffffffu, bit);
                peers &= bit ? m : ~m;
            }
#endif
            if (!valid) peers = 0;
            const u32 before = peers & lt;
            u32 wbase = 0;
            if (valid && before == 0) {
                wbase = whist16[warp * kParts + b];
                whist16[warp * kParts + b] = (unsigned short)(wbase + __popc(peers));
            }
            __syncwarp();
            wbase = __shfl_sync(0xffffffffu, wbase, (__ffs(peers) - 1) & 31);
            lrank[k] = (unsigned short)(wbase + __popc(before));
            bin[k] = valid ? b : 0xFFFFFFFFu;
        }
        __syncthreads();

        // (3) per partition: the tile's count (published at once), every warp's start inside the run
        u32 cnt[WPT][2];
        {
            u32 sum = 0;
#pragma unroll
            for (int j = 0; j < WPT; j++) {
                const u32 w = tid + j * THREADS;
                u32 run = 0;
#pragma unroll
                for (int ww = 0; ww < kWarps; ww++) { const u32 v = s.whist[ww][w]; s.whist[ww][w] = run; run += v; }
                cnt[j][0] = run & 0xFFFFu; cnt[j][1] = run >> 16;
                u32* lb = a.lookback + (u64)tile * kParts + 2 * w;
                const u32 fl = tile == 0 ? kLbInclusive : kLbPartial;
                st_relaxed_u32(lb, fl | cnt[j][0]);
                st_relaxed_u32(lb + 1, fl | cnt[j][1]);
                sum += cnt[j][0] + cnt[j][1];
            }
            // partition order for the staging area: word-major over threads (thread t's words are t, t + THREADS, ...),
            // any fixed order will do as long as (4) and (6) agree with it
            u32 total;
            u32 ex = block_excl_scan(sum, s.scan_part, &total);
#pragma unroll
            for (int j = 0; j < WPT; j++) {
                const u32 w = tid + j * THREADS;
                s.tile_excl[2 * w] = ex; ex += cnt[j][0];
                s.tile_excl[2 * w + 1] = ex; ex += cnt[j][1];
            }
            if (tid == 0) s.survivors = total;
        }
        __syncthreads();
        const u32 next_tile = s.next_tile;
        const u32 survivors = s.survivors;

        // (4) every record to its place in partition order
#pragma unroll
        for (int k = 0; k < ITEMS; k++) {
            if (bin[k] != 0xFFFFFFFFu) {
                const u32 j = warp * kWarpRecs + k * 32 + lane;
                const u32 b = bin[k];
                const u32 pos = s.tile_excl[b] + whist16[warp * kParts + b] + lrank[k];
                const u32 ty = s.ty[j], ph = s.ph[j];
                if (ty > kTypeDeleted || ph > kPhMax || (hi[k] | lo[k]) == 0) flags |= kFlagBadEvent;   // the run fails anyway
                del_surv += ty == kTypeDeleted;
                s.stage[pos * 3] = hi[k];
                s.stage[pos * 3 + 1] = lo[k];
                s.stage[pos * 3 + 2] = ((u64)uid_hash24(hi[k], lo[k]) << 38) | ((tile_base + j) << 8) | ((u64)(ty & 3u) << 6) | (eff_phase(ph) & 63u);
                s.sbin[pos] = (unsigned short)b;
            }
        }
        load_tile(next_tile);        // in flight during (5) and (6)

        // (5) decoupled look-back for this thread's partitions
#pragma unroll
        for (int j = 0; j < WPT; j++) {
            const u32 w = tid + j * THREADS;
            u32 excl[2] = {0, 0};
            if (tile > 0) {
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const u32* col = a.lookback + 2 * w + h;
                    int t = (int)tile - 1;
                    bool done = false;
                    while (!done) {
                        u32 v[4];
#pragma unroll
                        for (int q = 0; q < 4; q++) v[q] = t - q >= 0 ? ld_relaxed_u32(col + (u64)(t - q) * kParts) : kLbInclusive;
                        int used = 0;
#pragma unroll
                        for (int q = 0; q < 4; q++) {
                            if (done || (v[q] >> 30) == 0) break;
                            excl[h] += v[q] & kLbValue;
                            used++;
                            if ((v[q] >> 30) == 2) done = true;
                        }
                        t -= used;
                        if (!done && used == 0) __nanosleep(40);
                    }
                    st_relaxed_u32(a.lookback + (u64)tile * kParts + 2 * w + h, kLbInclusive | (excl[h] + cnt[j][h]));
                }
            }
            s.dst_off[2 * w] = pbase[j][0] + excl[0] - s.tile_excl[2 * w];
            s.dst_off[2 * w + 1] = pbase[j][1] + excl[1] - s.tile_excl[2 * w + 1];
        }
        __syncthreads();

        // (6) runs out to global memory: a (tile, partition) run leaves as one contiguous piece
#if PTE_SCATTER_WORDWISE
#pragma unroll 4
        for (u32 w = tid; w < survivors * 3; w += THREADS) {
            const u32 rec = w / 3;
            __stcs(a.out + (u64)(s.dst_off[s.sbin[rec]] + rec) * 3 + (w - rec * 3), s.stage[w]);
        }
#else
        // a lane per record (three 8-byte stores 24 bytes apart; neighbouring lanes fill the gaps, L2 merges the sectors):
        // a third of the index arithmetic of the word-by-word form
#pragma unroll 2
        for (u32 rec = tid; rec < survivors; rec += THREADS) {
            u64* d = a.out + (u64)(s.dst_off[s.sbin[rec]] + rec) * 3;
            const u64 w0 = s.stage[rec * 3], w1 = s.stage[rec * 3 + 1], w2 = s.stage[rec * 3 + 2];
            __stcs(d, w0); __stcs(d + 1, w1); __stcs(d + 2, w2);
        }
#endif
        tile = next_tile;
        __syncthreads();
    }
    del_surv = __reduce_add_sync(0xffffffffu, del_surv);
    flags = __reduce_or_sync(0xffffffffu, flags);
    if (lane == 0) {
        if (del_surv) atomicAdd(&a.ctr->n_del_survived, (u64)del_surv);
        if (flags) atomicOr(&a.ctr->flags, flags);
    }
}

// ---- bulk async copy + mbarrier (PTX; SASS: UBLKCP / SYNCS) ---------------------------------------------------
__device__ __forceinline__ u32 smem_addr(const void* p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(u64* bar, u32 count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(u64* bar, u32 bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_copy_g2s(void* dst, const void* src, u32 bytes, u64* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_addr(dst)), "l"(src), "r"(bytes), "r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(u64* bar, u32 parity) {
    u32 done;
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}\n" : "=r"(done) : "r"(smem_addr(bar)), "r"(parity) : "memory");
    return done != 0;
}
// Waiting warps must not eat the issue slots of the warps they wait for: back off between polls.
__device__ __forceinline__ void mbar_wait(u64* bar, u32 parity, u32 sleep_ns) {
    while (!mbar_try_wait(bar, parity)) if (sleep_ns) __nanosleep(sleep_ns);
}

// ---- dictionary slice with fingerprint tags ---------------------------------------------------------------
// Inside a partition the dictionary slice is used as lines of eight 16-byte slots (one 128-byte line).  A uid's
// home line comes from its hash; it is stored in the first slot of the line (then of the following lines) that was
// empty when it arrived, and slots never change once filled.  The CTA keeps one TAG byte per slot in shared
// memory: 0 = not known to be filled, else an 8-bit fingerprint of the uid in the slot.  A lookup reads the
// line's eight tags (one shared-memory load), loads only the slot whose tag matches and compares all 128 bits:
// one L2 round trip per event instead of a probe run of dependent loads.  Tags are a cache, the slots are the
// truth: a stale 0 tag (another CTA replaying a different segment of the same partition filled the slot) makes
// the 128-bit CAS fail, which returns the slot's content and repairs the tag.
struct UidHash { u32 line, fp; };
__device__ __forceinline__ UidHash uid_hash_of_bits(u32 h, u32 line_mask) {
    UidHash r;
    r.line = h & line_mask;
    r.fp = h >> 16;
    r.fp += r.fp == 0;
    return r;
}
__device__ __forceinline__ UidHash uid_hash(u64 hi, u64 lo, u32 line_mask) {
    const u32 h = uid_hash24(hi, lo);
    UidHash r;
    r.line = h & line_mask;            // at most 13 bits (slices of up to 65536 slots)
    r.fp = h >> 16;
    r.fp += r.fp == 0;
    return r;
}
__device__ __forceinline__ u64 bytes_equal(u64 t8, u32 b) {      // 0x80 in every byte of t8 equal to b (lowest hit exact, higher ones may be spurious)
    const u64 rep = 0x0101010101010101ull;
    const u64 x = t8 ^ (rep * b);
    return (x - rep) & ~x & (rep << 7);
}
// full find-or-insert; `tags` is this CTA's tag array of the slice
__device__ u32 intern_tagged(Slot* __restrict__ slice, u8* tags, u32 line_mask, u64 hi, u64 lo, u32& inserted,
                             u32* __restrict__ flags, u64 pol) {
    UidHash uh = uid_hash(hi, lo, line_mask);
    u32 line = uh.line;
    for (u32 walked = 0; walked <= line_mask;) {
        const u64 t8 = *reinterpret_cast<volatile u64*>(tags + line * 8);
        u64 m = bytes_equal(t8, uh.fp);
        while (m) {
            const u32 c = (u32)(__ffsll((long long)m) - 1) >> 3;
            m &= m - 1;
            const ulonglong2 v = ld_slot_cg(slice + line * 8 + c, pol);
            if (v.x == hi && v.y == lo) return line * 8 + c;
        }
        const u64 e8 = bytes_equal(t8, 0);
        if (e8) {
            const u32 e = (u32)(__ffsll((long long)e8) - 1) >> 3;
            const ulonglong2 old = cas128(slice + line * 8 + e, 0, 0, hi, lo);
            if (old.x == 0 && old.y == 0) { tags[line * 8 + e] = (u8)uh.fp; inserted++; return line * 8 + e; }
            tags[line * 8 + e] = (u8)uid_hash(old.x, old.y, line_mask).fp;       // filled by somebody else: repair the tag
            if (old.x == hi && old.y == lo) return line * 8 + e;
            continue;                                                            // same line again, with the new tag
        }
        line = (line + 1) & line_mask;
        walked++;
    }
    atomicOr(flags, kFlagTableFull);
    return 0;
}

// ---- replay of the partitions ---------------------------------------------------------------------------
// One CTA of 1024 threads per segment, warp-specialised, tiles of 2048 records in a ring of three stages:
//   * the tile's bytes arrive by one bulk async copy (cp.async.bulk, completion on an mbarrier);
//   * sixteen PROBER warps intern the uids in the partition's dictionary slice (L2 hits; find-or-insert with the
//     128-bit CAS) and deal every record, keeping arrival order, to the replayer that owns its slot (slot mod 16):
//     ballots over the four queue bits, per-warp counts, one scan over the 16 x 16 counts;
//   * sixteen REPLAYER warps each walk their queue of the tile in order, 32 records per step, against the state
//     bytes of their own pods in shared memory.  A step whose lanes all have different slots (the usual case,
//     found with a mark byte per slot) is one load and one store per lane; otherwise MATCH.ANY gives the exact
//     peers and the lanes of a slot chain their states in lane order.
// A lone warp runs such a dependent step in 400+ cycles whatever the instruction count, which is why the tile's
// pods are spread over sixteen of them; the probers run a stage or two ahead, so the dictionary's latency
// never sits in a replayer's chain.  When all replayers are done with a stage, one of them refills it.
static constexpr int kRpThreads = 1024, kRpWarps = kRpThreads / 32, kRpTile = 2048;
static constexpr int kRpReplayers = 16, kRpProberWarps = kRpWarps - kRpReplayers, kRpProberThreads = kRpProberWarps * 32;
static constexpr int kRpPerProber = kRpTile / kRpProberThreads;      // records per prober thread and tile
static constexpr int kRpWarpRecs = kRpTile / kRpProberWarps;         // consecutive records one prober warp deals
static constexpr int kRpQBits = 4;
static constexpr int kRpBufRecs = kRpTile + 2;        // a tile is copied from an even record index: 16-byte aligned
static constexpr int kRpStages = 3;
static constexpr int kRpMarks = 1024;                 // mark bytes per replayer
static_assert(kRpReplayers == 1 << kRpQBits && kRpPerProber * kRpProberThreads == kRpTile, "replay geometry");

struct ReplayArgs {
    const u64* rec;                 // partition-ordered records, 3 words each
    const Segment* segs; const u32* n_segs; u32* ticket;
    u32 first_seg;                  // index of segs[0] in the global numbering (0; kMaxSegments for the sample)
    Slot* table; u32 slice_slots;   // dictionary: kParts slices of slice_slots slots
    u32* bitmap;
    Counters* ctr;
    u8* seg_export;                 // [segment][slice_slots] final state bytes of segments that are followed by another
    PendingHead* pending; u32 pending_cap;
    int count_only;
};

// dynamic shared memory: [kRpStages][kRpBufRecs*3] u64 | state[slice_slots] | tags[slice_slots] | the fixed part below
struct ReplaySmemFixed {
    unsigned short sslot[kRpStages][kRpTile];     // dictionary slot of every record of the tile
    unsigned short queue[kRpStages][kRpTile];     // per dealing warp (kRpWarpRecs entries each): its records grouped by replayer, arrival order inside
    unsigned short wstart[kRpStages][kRpProberWarps][kRpReplayers + 1];   // where each replayer's group starts in a dealing warp's part
    u32 wcnt[kRpProberWarps][kRpReplayers];       // dealing warp's running count per replayer
    u8 mark[kRpReplayers][kRpMarks];              // per replayer: lane that last claimed a slot in the current step
    u64 bar_tile[kRpStages];                      // bulk copy of the stage has landed
    u64 bar_slots[kRpStages];                     // all prober warps are done with the stage
    u64 bar_done[kRpStages];                      // all replayer warps are done with the stage
    u32 seg;
};

struct ReplayCtx {
    const ReplayArgs* a; u32 seg_id; bool invert;
    u32 n_emit, n_seen;
};
// one event against the state `pred` of its pod
__device__ __forceinline__ void replay_event(ReplayCtx& c, u32 pred, u64 meta, u32 slot) {
    const u32 tp = (u32)meta & 0xFFu;
    const ReplayArgs& a = *c.a;
    if (pred == 0xFFu) {
        // first event of this uid in a later segment: the predecessor lives in an earlier segment
        if (!a.count_only) {
            const u32 at = atomicAdd(&a.ctr->n_pending, 1u);
            if (at < a.pending_cap) { PendingHead p; p.meta = meta; p.seg = c.seg_id; p.slot = slot; a.pending[at] = p; }
            else atomicOr(&a.ctr->flags, kFlagPendingFull);
        }
    } else {
        const bool e = emits(pred, tp);
        c.n_emit += e; c.n_seen++;
        if (!a.count_only && e != c.invert) flip_bit(a.bitmap, (u32)(meta >> 8) & 0x3FFFFFFFu);
    }
}

__device__ __forceinline__ void mbar_arrive(u64* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}
// the replayer that owns a slot: slots of a line fill from its low end, so the line index is folded in
__device__ __forceinline__ u32 slot_owner(u32 slot) { return (slot ^ (slot >> 3)) & (kRpReplayers - 1); }
__device__ __forceinline__ void prober_barrier() { asm volatile("bar.sync 1, %0;" ::"n"(kRpProberThreads) : "memory"); }

#ifdef PTE_EXP_TRACE
__device__ unsigned long long g_rp_trace[64 * 8];
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define RP_TR(slot, cond) if (blockIdx.x == 0 && (cond) && t >= 200 && t < 264 && !a.count_only && seg_iter == 1) g_rp_trace[(t - 200) * 8 + (slot)] = gtime();
#define RP_TRMAX(slot, cond) if (blockIdx.x == 0 && (cond) && t >= 200 && t < 264 && !a.count_only && seg_iter == 1) atomicMax(&g_rp_trace[(t - 200) * 8 + (slot)], gtime());
#else
#define RP_TR(slot, cond)
#define RP_TRMAX(slot, cond)
#endif
__global__ void __launch_bounds__(kRpThreads, 1) part_replay_kernel(ReplayArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    u64* buf = reinterpret_cast<u64*>(smem_raw);
    u8* state = smem_raw + kRpStages * kRpBufRecs * 24;
    u8* tags = state + ((a.slice_slots + 15) & ~15u);
    ReplaySmemFixed& f = *reinterpret_cast<ReplaySmemFixed*>(tags + ((a.slice_slots + 15) & ~15u));
    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const u32 lt = lanemask_lt();
    const u32 NS = a.slice_slots, line_mask = NS / 8 - 1;
    const u64 pol = l2_policy_evict_last();
    ReplayCtx cx;
    cx.a = &a; cx.invert = a.ctr->invert != 0; cx.n_emit = 0; cx.n_seen = 0; cx.seg_id = 0;
    u32 inserted = 0, seg_iter = 0;
    u32 par_a = 0, par_b = 0;      // phase parities, one bit per stage: probers (tile), replayers (slots, done)

    if (tid == 0)
        for (int b = 0; b < kRpStages; b++) {
            mbar_init(&f.bar_tile[b], 1); mbar_init(&f.bar_slots[b], kRpProberWarps); mbar_init(&f.bar_done[b], kRpReplayers);
        }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
    const u32 n_segs = *a.n_segs;

    for (;;) {
        if (tid == 0) f.seg = atomicAdd(a.ticket, 1u);
        __syncthreads();
        const u32 seg_local = f.seg;
        if (seg_local >= n_segs) break;
        const Segment sg = a.segs[seg_local];
        seg_iter++;
        cx.seg_id = a.first_seg + seg_local;
        const bool later = (sg.ordinal & 0x7FFFFFFFu) != 0, has_next = (sg.ordinal >> 31) != 0;
        const u32 n_rec = sg.end - sg.beg;
        const u32 n_tiles = (n_rec + kRpTile - 1) / kRpTile;
        Slot* slice = a.table + (u64)sg.part * NS;
        // state bytes: "no live phase" for a first segment, "not seen here" otherwise
        {
            const u32 fill = later ? 0xFFFFFFFFu : 0u;
            for (u32 k = tid; k < (NS + 3) / 4; k += kRpThreads) {
                reinterpret_cast<u32*>(state)[k] = fill;
                reinterpret_cast<u32*>(tags)[k] = 0;
            }
        }
        auto issue = [&](u32 t) {        // one thread: bulk copy of tile t into stage t % kRpStages
            const u32 first = sg.beg + t * kRpTile, even = first & ~1u;
            u64* bar = &f.bar_tile[t % kRpStages];
            mbar_expect_tx(bar, kRpBufRecs * 24);
            bulk_copy_g2s(buf + (t % kRpStages) * kRpBufRecs * 3, a.rec + (u64)even * 3, kRpBufRecs * 24, bar);
        };
        __syncthreads();                 // state initialised; everyone is done with the previous segment's stages
        if (tid == 0) for (u32 t = 0; t < kRpStages && t < n_tiles; t++) issue(t);

        if (warp >= kRpReplayers) {
            // ---- probers: records -> dictionary slots -> the owning replayer's queue
            const u32 pw = warp - kRpReplayers;
            for (u32 t = 0; t < n_tiles; t++) {
                const u32 sidx = t % kRpStages;
                const u32 first = sg.beg + t * kRpTile;
                const u32 valid = min((u32)kRpTile, sg.end - first);
                const u64* tb = buf + sidx * kRpBufRecs * 3 + (first & 1u) * 3;
                if (lane < kRpReplayers) f.wcnt[pw][lane] = 0;
                __syncwarp();
                RP_TR(0, pw == 0 && lane == 0)
                mbar_wait(&f.bar_tile[sidx], (par_a >> sidx) & 1u, 128);
                RP_TR(1, pw == 0 && lane == 0)
                par_a ^= 1u << sidx;
                // this warp's records are consecutive.  Lookups run in warp-wide ROUNDS: in every round each unresolved
                // record of a thread loads its next candidate slot (the next tag of its line that carries its fingerprint),
                // all loads of the round are in flight together, then they are compared.  Most records are done after the
                // first round; a fingerprint shared inside a line or a uid displaced to the next line costs one more
                // round for the whole warp -- not a dependent chain of L2 round trips per record.
                u32 r[kRpPerProber], slot_of[kRpPerProber], line[kRpPerProber], cm[kRpPerProber];
                u64 hi[kRpPerProber], lo[kRpPerProber];
                bool ok[kRpPerProber];
                u32 pend = 0, walked = 0;
                auto tag_masks = [&](u32 ln, u32 fp) -> u32 {      // bits 0..7 slots tagged fp, bits 8..15 slots tagged empty, 24..31 fp
                    const u64 t8 = *reinterpret_cast<volatile u64*>(tags + ln * 8);
                    const u64 sel = 0x0102040810204080ull;
                    const u32 mm = (u32)((((bytes_equal(t8, fp) >> 7) & 0x0101010101010101ull) * sel) >> 56);
                    const u32 me = (u32)((((bytes_equal(t8, 0) >> 7) & 0x0101010101010101ull) * sel) >> 56);
                    return mm | (me << 8) | (fp << 24);
                };
#pragma unroll
                for (int k = 0; k < kRpPerProber; k++) {
                    r[k] = pw * kRpWarpRecs + k * 32 + lane;
                    ok[k] = r[k] < valid;
                    hi[k] = ok[k] ? tb[r[k] * 3] : 0; lo[k] = ok[k] ? tb[r[k] * 3 + 1] : 0;
                    const UidHash uh = uid_hash_of_bits(ok[k] ? (u32)(tb[r[k] * 3 + 2] >> 38) & 0xFFFFFFu : 0u, line_mask);
                    line[k] = uh.line;
                    cm[k] = ok[k] ? tag_masks(uh.line, uh.fp) : 0u;
                    slot_of[k] = 0;
                    pend |= ok[k] ? 1u << k : 0u;
                }
                while (__any_sync(0xffffffffu, pend != 0)) {
                    u32 cand[kRpPerProber];
                    ulonglong2 pv[kRpPerProber];
#pragma unroll
                    for (int k = 0; k < kRpPerProber; k++) {
                        cand[k] = 0xFFFFFFFFu;
                        pv[k] = make_ulonglong2(0, 0);
                        if (((pend >> k) & 1u) && (cm[k] & 0xFFu)) {
                            // bytes_equal may flag a byte above a true hit spuriously: a wasted load, never a wrong answer
                            const u32 c = __ffs(cm[k] & 0xFFu) - 1;
                            cm[k] &= ~(1u << c);
                            cand[k] = line[k] * 8 + c;
                            pv[k] = ld_slot_cg(slice + cand[k], pol);
                        }
                    }
#pragma unroll
                    for (int k = 0; k < kRpPerProber; k++) {
                        if (!((pend >> k) & 1u)) continue;
                        if (cand[k] != 0xFFFFFFFFu) {
                            if (pv[k].x == hi[k] && pv[k].y == lo[k]) { slot_of[k] = cand[k]; pend &= ~(1u << k); }
                        } else if ((cm[k] >> 8) & 0xFFu) {
                            // no tagged candidate left and the line has a slot this CTA believes empty: a new uid (or a tag
                            // that is stale because another CTA filled the slot) -- the full find-or-insert sorts it out
                            slot_of[k] = intern_tagged(slice, tags, line_mask, hi[k], lo[k], inserted, &a.ctr->flags, pol);
                            pend &= ~(1u << k);
                        } else if (++walked > 4 * (line_mask + 1)) {
                            // every line of the slice is full and none holds the uid: more pods than the slice has room for
                            atomicOr(&a.ctr->flags, kFlagTableFull);       // the run is repeated on the sort path
                            pend &= ~(1u << k);
                        } else {
                            line[k] = (line[k] + 1) & line_mask;       // the line is full: the uid may have been displaced to the next
                            cm[k] = tag_masks(line[k], cm[k] >> 24);
                        }
                    }
                }
                u32 q[kRpPerProber];
                unsigned short lrank[kRpPerProber];
#pragma unroll
                for (int k = 0; k < kRpPerProber; k++) {
                    const u32 slot = slot_of[k];
                    if (ok[k]) f.sslot[sidx][r[k]] = (unsigned short)slot;
                    const u32 qq = slot_owner(slot);
                    u32 peers = __ballot_sync(0xffffffffu, ok[k]);
#pragma unroll
                    for (int b = 0; b < kRpQBits; b++) {
                        const bool bit = (qq >> b) & 1u;
                        const u32 m = __ballot_sync(0xffffffffu, bit);
                        peers &= bit ? m : ~m;
                    }
                    if (!ok[k]) peers = 0;
                    const u32 before = peers & lt;
                    u32 wb = 0;
                    if (ok[k] && before == 0) { wb = f.wcnt[pw][qq]; f.wcnt[pw][qq] = wb + __popc(peers); }
                    __syncwarp();
                    wb = __shfl_sync(0xffffffffu, wb, (__ffs(peers) - 1) & 31);
                    lrank[k] = (unsigned short)(wb + __popc(before));
                    q[k] = qq;
                }
                __syncwarp();
                // this warp's counts -> where each replayer's group starts inside the warp's part of the queue (no block
                // barrier: the replayers stitch the sixteen parts together themselves)
                {
                    const u32 c = lane < kRpReplayers ? f.wcnt[pw][lane] : 0u;
                    u32 ex = c;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) { const u32 x = __shfl_up_sync(0xffffffffu, ex, o); if (lane >= (u32)o) ex += x; }
                    if (lane <= kRpReplayers) f.wstart[sidx][pw][lane] = (unsigned short)(ex - c);    // lane 16: the warp's total
                }
                __syncwarp();
#pragma unroll
                for (int k = 0; k < kRpPerProber; k++)
                    if (ok[k]) f.queue[sidx][pw * kRpWarpRecs + f.wstart[sidx][pw][q[k]] + lrank[k]] = (unsigned short)r[k];
                __syncwarp();
                RP_TR(2, pw == 0 && lane == 0)
                RP_TRMAX(6, lane == 0)
                if (lane == 0) mbar_arrive(&f.bar_slots[sidx]);
            }
        } else {
            // ---- replayers: this warp's queue of the tile, in arrival order, 32 records per step
            u8* mark = f.mark[warp];
            for (u32 t = 0; t < n_tiles; t++) {
                const u32 sidx = t % kRpStages;
                const u32 first = sg.beg + t * kRpTile;
                const u64* tb = buf + sidx * kRpBufRecs * 3 + (first & 1u) * 3;
                mbar_wait(&f.bar_slots[sidx], (par_a >> sidx) & 1u, 32);      // the probers waited for the tile's bytes
                RP_TR(3, warp == 0 && lane == 0)
                par_a ^= 1u << sidx;
                const unsigned short* ss = f.sslot[sidx];
                const unsigned short* qu = f.queue[sidx];
                // this replayer's records sit in sixteen groups, one per dealing warp (lane l < 16 looks after warp l's)
                u32 g_beg = 0, g_len = 0;
                if (lane < kRpProberWarps) {
                    g_beg = f.wstart[sidx][lane][warp];
                    g_len = f.wstart[sidx][lane][warp + 1] - g_beg;
                    g_beg += lane * kRpWarpRecs;
                }
                u32 g_ex = g_len;
#pragma unroll
                for (int o = 1; o < 16; o <<= 1) { const u32 x = __shfl_up_sync(0xffffffffu, g_ex, o); if (lane >= (u32)o) g_ex += x; }
                const u32 qe = __shfl_sync(0xffffffffu, g_ex, kRpProberWarps - 1);       // records of this replayer in the tile
                g_ex -= g_len;                                                          // exclusive: first queue position of group l
                auto fetch = [&](u32 i, u32& s_out, u64& m_out) {                       // i-th record of this replayer's queue
                    // the group that holds position i: the last one that starts at or before i (binary search over lanes 0..15)
                    u32 g = 0;
#pragma unroll
                    for (int step = kRpProberWarps / 2; step; step >>= 1) {
                        const u32 probe = __shfl_sync(0xffffffffu, g_ex, g + step);
                        if (probe <= i) g += step;
                    }
                    const u32 gb = __shfl_sync(0xffffffffu, g_beg, g), ge = __shfl_sync(0xffffffffu, g_ex, g);
                    s_out = 0xFFFF0000u | lane; m_out = 0;
                    if (i < qe) { const u32 rec = qu[gb + (i - ge)]; s_out = ss[rec]; m_out = tb[rec * 3 + 2]; }
                };
                u32 s_next;
                u64 m_next;
                fetch(lane, s_next, m_next);
                for (u32 base = 0; base < qe; base += 32) {
                    const bool active = base + lane < qe;
                    const u32 s = s_next;
                    const u64 meta = m_next;
                    if (base + 32 < qe) fetch(base + 32 + lane, s_next, m_next);      // next step's operands, off the chain
                    const u32 tp = (u32)meta & 0xFFu;
                    const u32 h = (s >> kRpQBits) & (kRpMarks - 1);
                    if (active) mark[h] = (u8)lane;
                    __syncwarp();
                    u32 pred = 0;
                    bool clash = false;
                    if (active) { clash = mark[h] != lane; pred = state[s]; }
                    if (__any_sync(0xffffffffu, clash)) {
                        // some lanes may share a slot: exact peers, states chained in lane order
                        const u32 peers = __match_any_sync(0xffffffffu, s);
                        const u32 lower = peers & lt;
                        const u32 src = lower ? 31u - __clz(lower) : lane;
                        const u32 ptp = __shfl_sync(0xffffffffu, tp, src);
                        if (lower) pred = state_after(ptp);
                        if (active && (peers >> lane) == 1u) state[s] = (u8)state_after(tp);      // last of its slot in this step
                    } else if (active) {
                        state[s] = (u8)state_after(tp);
                    }
                    if (active) replay_event(cx, pred, meta, s);
                    __syncwarp();                                             // this step's state before the next step's loads
                }
                RP_TR(4, warp == 0 && lane == 0)
                RP_TRMAX(7, lane == 0)
                if (lane == 0) mbar_arrive(&f.bar_done[sidx]);
                if (warp == 0 && t + kRpStages < n_tiles) {
                    // the stage is free once every replayer is done with it: refill
                    mbar_wait(&f.bar_done[sidx], (par_b >> sidx) & 1u, 32);
                    RP_TR(5, lane == 0)
                    if (lane == 0) issue(t + kRpStages);
                }
                if (warp == 0) par_b ^= 1u << sidx;
            }
        }
        __syncthreads();                 // the state bytes are final
        if (has_next && !a.count_only) {
            u32* dst = reinterpret_cast<u32*>(a.seg_export + (u64)cx.seg_id * NS);
            for (u32 k = tid; k < NS / 4; k += kRpThreads) dst[k] = reinterpret_cast<const u32*>(state)[k];
        }
    }
    inserted = __reduce_add_sync(0xffffffffu, inserted);
    if (lane == 0 && inserted) atomicAdd(&a.ctr->n_pods, inserted);
    if (a.count_only && warp < kRpReplayers) {
        const u32 n_emit = __reduce_add_sync(0xffffffffu, cx.n_emit);
        const u32 n_seen = __reduce_add_sync(0xffffffffu, cx.n_seen);
        if (lane == 0 && n_seen) { atomicAdd(&a.ctr->sample_emit, n_emit); atomicAdd(&a.ctr->sample_total, n_seen); }
    }
}

// First events of a uid inside a later segment: find the state the nearest earlier segment of the same
// partition left for that slot (0xFF = that segment never saw the uid), else "never seen".
__global__ void __launch_bounds__(256)
pending_fixup_kernel(const PendingHead* __restrict__ pending, u32 pending_cap, const Segment* __restrict__ segs,
                     const u8* __restrict__ seg_export, u32 slice_slots, u32* __restrict__ bitmap, Counters* __restrict__ ctr) {
    const u32 n = min(ctr->n_pending, pending_cap);
    const bool invert = ctr->invert != 0;
    for (u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const PendingHead p = pending[i];
        const u32 ord = segs[p.seg].ordinal & 0x7FFFFFFFu;
        u32 pred = 0;
        for (u32 back = 1; back <= ord; back++) {
            const u32 st = seg_export[(u64)(p.seg - back) * slice_slots + p.slot];
            if (st != 0xFFu) { pred = st; break; }
        }
        if (emits(pred, (u32)p.meta & 0xFFu) != invert) flip_bit(bitmap, (u32)(p.meta >> 8) & 0x3FFFFFFFu);
    }
}

// Interned key of each emitted record after a partitioned run: partition * slice_slots + slot.
__global__ void __launch_bounds__(256)
key_lookup_part_kernel(const u32* __restrict__ out_idx, u64 n, const u64* __restrict__ uid_hi, const u64* __restrict__ uid_lo,
                       const Slot* __restrict__ table, u32 slice_slots, u64* __restrict__ out_key) {
    const u32 line_mask = slice_slots / 8 - 1;
    for (u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (u64)gridDim.x * blockDim.x) {
        const u32 i = out_idx[j];
        const u64 hi = uid_hi[i], lo = uid_lo[i];
        const u32 part = part_of(lo);
        const Slot* slice = table + (u64)part * slice_slots;
        u32 sl = uid_hash(hi, lo, line_mask).line * 8;
        u64 key = ~0ull;
        for (u32 p = 0; p < slice_slots; ++p) {           // slots in probe order: the uid sits before the first empty one
            const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(slice + sl);
            if (v.x == hi && v.y == lo) { key = (u64)part * slice_slots + sl; break; }
            if ((v.x | v.y) == 0) break;
            sl = (sl + 1) & (slice_slots - 1);
        }
        out_key[j] = key;
    }
}
