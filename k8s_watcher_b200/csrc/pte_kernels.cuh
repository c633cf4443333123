// Device kernels of the replay path other than the sort (pte_radix.cuh) and the JSON scanner (pte_json.cuh):
//   filter_flag_kernel    the reference's two predicates (pod_watcher.py:204-212, :226-229) + rv order check -> keep bitmap
//   intern_pack_kernel    uid interning (exact, 128-bit CAS dictionary) + record packing + digit histograms
//   diff_sample/bitmap_init/diff_flag kernels
//                         segmented adjacent-diff on the key-sorted records -> emit bitmap,
//                         optionally against the per-pod state carried from earlier batches of a stream
//   bitmap_count/tile_scan/assemble_bitmap kernels
//                         scan + compaction of the bitmap in replay order -> payload columns (pod_watcher.py:232-236)
//   key_lookup_kernel     interned key of every emitted record, on demand
//   payload_* kernels     request bodies as `requests` would send them (clusterapi_client.py:36)
//   rv_* / permute_log / map_idx kernels   pre-sort of logs that are not in resourceVersion order
//   digest_kernel         order-sensitive digest for large-scale parity
//   gen_* kernels         synthetic replay logs (spec in include/pte.h)
#pragma once
#include "pte_common.cuh"
#include "pte_radix.cuh"

struct __align__(16) Slot { u64 hi, lo; };

enum : u32 { kFlagUnordered = 1u, kFlagBadEvent = 2u, kFlagTableFull = 4u };

// Device-resident counters of one run (zeroed by the host before each pte_run).
struct Counters {
    u64 n_survived;
    u64 n_emitted;
    u64 n_tombstones;      // emitted DELETED (T-trace)
    u64 n_del_survived;    // DELETED among survivors (R-trace tombstones)
    u64 digest;
    u32 n_pods;
    u32 flags;
    u32 ticket[4];         // tile tickets: [0] spare, [1] radix pass (reset per pass), [2] partition scatter, [3] partition replay
    u32 sample_emit, sample_total;   // emit fraction of a sample of the sorted records
    u32 invert;            // 1: the diff stage clears the bits of suppressed records instead of setting emitted ones
    u32 n_pending;         // partitioned runs: first events of a uid in a later segment, resolved by pending_fixup_kernel
};

__device__ __forceinline__ ulonglong2 cas128(Slot* addr, u64 cmp_hi, u64 cmp_lo, u64 val_hi, u64 val_lo) {
    ulonglong2 old;
    asm volatile(
        "{\n\t"
        ".reg .b128 c, v, o;\n\t"
        "mov.b128 c, {%2, %3};\n\t"
        "mov.b128 v, {%4, %5};\n\t"
        "atom.global.relaxed.gpu.cas.b128 o, [%6], c, v;\n\t"
        "mov.b128 {%0, %1}, o;\n\t"
        "}\n"
        : "=l"(old.x), "=l"(old.y)
        : "l"(cmp_hi), "l"(cmp_lo), "l"(val_hi), "l"(val_lo), "l"(addr)
        : "memory");
    return old;
}

static constexpr int kMaxProbe = 2048;

// Dictionary slots are read with an L2 evict_last hint, the log streams through with evict_first (ld/st .cs).
// Measured (profiles/README.md, r01_gather/): the hint changes nothing -- a random probe over a table of twice
// the L2 hits 26 % of the time with or without it, a miss always brings in the whole 128-byte line, and the
// kernel is bound by the rate of random lines (46 G/s of the 54 G/s a bare gather reaches), not by bytes.
// PTE_SLOT_LD picks the load for tuning builds: 0 evict_last hint (default), 1 L2::64B (half the DRAM bytes per
// miss, but 26.0 instead of 23.7 ms: the next slot of a probe run misses more often), 2 both, 3 plain.
#ifndef PTE_SLOT_LD
#define PTE_SLOT_LD 0
#endif
__device__ __forceinline__ u64 l2_policy_evict_last() {
    u64 pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ ulonglong2 ld_slot(const Slot* p, u64 pol) {
    ulonglong2 v;
#if PTE_SLOT_LD == 1
    asm volatile("ld.global.L2::64B.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p) : "memory");
#elif PTE_SLOT_LD == 2
    asm volatile("ld.global.L2::cache_hint.L2::64B.v2.u64 {%0, %1}, [%2], %3;" : "=l"(v.x), "=l"(v.y) : "l"(p), "l"(pol) : "memory");
#elif PTE_SLOT_LD == 3
    asm volatile("ld.global.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p) : "memory");
#else
    asm volatile("ld.global.L2::cache_hint.v2.u64 {%0, %1}, [%2], %3;" : "=l"(v.x), "=l"(v.y) : "l"(p), "l"(pol) : "memory");
#endif
    return v;
}

// uid -> slot of an open-addressing table (linear probing, 16-byte slots, 0:0 = empty).
// Exact: two uids share a slot only if all 128 bits are equal.
// A plain 16-byte load may see a slot as empty when it is not (stale line), or half-written
// (0:0 -> hi:lo is the only transition a slot ever makes).  Either shows up as a zero half,
// and then the 128-bit CAS -- which returns the true content -- decides.
// `v` is the content of slot `sl` as loaded by the caller (so that a thread can have the first
// probe of all its events in flight before it resolves any of them).
// p-th slot of the probe sequence that starts at `home`.  The sequence walks the eight slots of home's
// 128-byte line cyclically before it moves on to the next line (same offset), because a probe that misses
// L2 brings in the whole line (profiles/r01_gather/): a probe run then leaves its first line only when the
// line is full.  PTE_PROBE_LINE=0 is plain linear probing.
#ifndef PTE_PROBE_LINE
#define PTE_PROBE_LINE 1
#endif
__device__ __forceinline__ u32 probe_slot(u32 home, u32 p, u32 mask) {
#if PTE_PROBE_LINE == 2      // experiment: runs confined to a 64-byte half line (with PTE_SLOT_LD=1: 64-byte L2 fetches)
    return ((((home >> 2) + (p >> 2)) << 2) | ((home + p) & 3u)) & mask;
#elif PTE_PROBE_LINE
    return ((((home >> 3) + (p >> 3)) << 3) | ((home + p) & 7u)) & mask;
#else
    return (home + p) & mask;
#endif
}
// the same sequence, stepped: slot p from slot p-1 (needs no copy of `home` in a register)
__device__ __forceinline__ u32 probe_next(u32 sl, u32 p, u32 mask) {
#if PTE_PROBE_LINE == 2
    return (((sl & ~3u) + ((p & 3u) == 0 ? 4u : 0u)) | ((sl + 1) & 3u)) & mask;
#elif PTE_PROBE_LINE
    return (((sl & ~7u) + ((p & 7u) == 0 ? 8u : 0u)) | ((sl + 1) & 7u)) & mask;
#else
    return (sl + 1) & mask;
#endif
}

// L2-only flavour of ld_slot (ld.global.cg): for kernels in which one SM probes the same small dictionary slice over
// and over.  An L1-cached probe keeps seeing a slot as empty long after another thread filled it, and every such
// probe falls through to the 128-bit CAS (measured: 40 ms instead of 1 ms per 10^9 probes).
__device__ __forceinline__ ulonglong2 ld_slot_cg(const Slot* p, u64 pol) {
    ulonglong2 v;
    asm volatile("ld.global.cg.L2::cache_hint.v2.u64 {%0, %1}, [%2], %3;" : "=l"(v.x), "=l"(v.y) : "l"(p), "l"(pol) : "memory");
    return v;
}

// p0: how many slots of the run have been looked at already (sl is slot p0 of the run, v its content)
template <bool CG = false>
__device__ __forceinline__ u32 intern_uid(Slot* __restrict__ table, u32 mask, u64 hi, u64 lo, u32 sl, ulonglong2 v,
                                          u32& inserted, u32* __restrict__ flags, u64 pol, int p0 = 0) {
    for (int p = p0; p < kMaxProbe; ++p) {
        if (v.x == 0 || v.y == 0) {
            ulonglong2 old = cas128(table + sl, 0, 0, hi, lo);
            if (old.x == 0 && old.y == 0) { inserted++; return sl; }
            if (old.x == hi && old.y == lo) return sl;
        } else if (v.x == hi && v.y == lo) {
            return sl;
        }
        sl = probe_next(sl, p + 1, mask);
        v = CG ? ld_slot_cg(table + sl, pol) : ld_slot(table + sl, pol);
    }
    atomicOr(flags, kFlagTableFull);
    return 0;
}
__device__ __forceinline__ u32 uid_home_slot(u64 hi, u64 lo, u32 mask) {
    return (u32)mix64(hi ^ (lo * 0x9E3779B97F4A7C15ull)) & mask;
}

struct InternArgs {
    const u64* uid_hi; const u64* uid_lo; const u64* rv;
    const u32* ns; const u8* phase; const u8* type;
    u64 n;
    Slot* table; u32 table_mask;
    RecLayout L;
    u64* rec_out;
    const u32* ns_bitmap; u32 ns_words; int ns_filter_on; int critical_only;
    Counters* ctr;
    const u32* keep;       // FILTERED: keep bitmap written by filter_flag_kernel
    const u32* block_off;  // FILTERED: records before each 8192-event block (tile_scan_kernel)
    u32* hist;         // digit counts of the key field for the radix passes that follow
    int key_passes;    // number of 8-bit digits in the key field (<= 4)
};

static constexpr int kAsmTile = 256 * 32;       // positions per bitmap tile (256 words)
#ifndef PTE_ASM_ROWS
#define PTE_ASM_ROWS 4
#endif
static constexpr int kAsmRows = PTE_ASM_ROWS;   // rows a thread of assemble_bitmap_kernel keeps in flight
static constexpr int kInternThreads = 256;
static constexpr int kInternPerThread = 4;
static constexpr int kInternTile = kInternThreads * kInternPerThread;

// The reference's two predicates, on integer codes.
template <typename Args>
__device__ __forceinline__ bool passes_filters(const Args& a, u32 ty, u32 ph, u32 ns) {
    // should_process_event, pod_watcher.py:209
    if (a.critical_only && ty != kTypeDeleted && ph != kPhNoStatus && ph != kPhFailed && ph != kPhSucceeded) return false;
    // namespace allow-list, pod_watcher.py:227
    if (a.ns_filter_on) {
        const u32 w = ns >> 5;
        return w < a.ns_words && ((__ldg(a.ns_bitmap + w) >> (ns & 31)) & 1u);
    }
    return true;
}

// uid interning + record packing.  Thread t of a 1024-event tile owns events 4t..4t+3 (16-byte vector
// loads).  The record's position field is always the arrival index of the event.
// FILTERED == false: no predicate is active, event i becomes record i (two 16-byte streaming stores per thread).
// FILTERED == true : filter_flag_kernel has already evaluated the reference's predicates into a keep
//                    bitmap, and bitmap_count/tile_scan have turned it into one output offset per
//                    8192-event block; survivors are compacted in arrival order behind that offset.  No
//                    ticket and no look-back: the blocks are independent.
// (Interning in several rounds over L2-sized slices of the dictionary was tried and is slower: probes
//  that hit L2 only halve the kernel, and every round has to re-read the uid columns.
//  Advancing the probe runs of a thread's four events together, so that second and third probes overlap, was tried
//  too -- the dictionary's load costs 2.6 ms of the kernel, profiles/r02_yardsticks/table_load_sweep.txt -- and is 2-4 x
//  slower: all lanes wait for the longest run of 128 events, and at 48 registers the extra state spills.)
#ifndef PTE_INTERN_BLOCKS
#define PTE_INTERN_BLOCKS 5      // measured on B200: 3/4/5/6 CTAs per SM -> 29.1/24.7/23.7/28.3 ms (6 spills too much)
#endif
template <bool FILTERED>
__global__ void __launch_bounds__(kInternThreads, FILTERED ? 4 : PTE_INTERN_BLOCKS) intern_pack_kernel(InternArgs a) {
    __shared__ u32 s_tot[8];
    __shared__ u32 s_hist[4 * kRadix];     // the sort's digit counts are taken here: no extra read of the records
    const u32 tid = threadIdx.x, lane = tid & 31;
    u32 inserted = 0, del_surv = 0, flags = 0;
    if (!FILTERED && blockIdx.x == 0 && tid == 0) a.ctr->n_survived = a.n;
    for (int k = tid; k < 4 * kRadix; k += kInternThreads) s_hist[k] = 0;
    __syncthreads();
    const u32 n_tiles = (u32)((a.n + kInternTile - 1) / kInternTile);
    const u64 pol = l2_policy_evict_last();
    constexpr u32 kSub = kAsmTile / kInternTile;          // 1024-event tiles per 8192-event block
    u32 run = 0, cur_block = ~0u;                         // FILTERED: records written so far in the current block

    // FILTERED walks whole 8192-event blocks (its tiles must come in order inside a block)
    const u32 step = FILTERED ? gridDim.x * kSub : gridDim.x;
    for (u32 tile0 = FILTERED ? blockIdx.x * kSub : blockIdx.x; tile0 < n_tiles; tile0 += step)
    for (u32 sub = 0; sub < (FILTERED ? kSub : 1u); sub++) {
        const u32 tile = tile0 + sub;
        if (tile >= n_tiles) break;
        const u64 i0 = (u64)tile * kInternTile + (u64)tid * kInternPerThread;

        u64 hi_[4], lo_[4];
        u32 ph[4], ty[4];
        {
            const ulonglong2* p = reinterpret_cast<const ulonglong2*>(a.uid_hi + i0);
            ulonglong2 x = __ldcs(p), y = __ldcs(p + 1);
            hi_[0] = x.x; hi_[1] = x.y; hi_[2] = y.x; hi_[3] = y.y;
            p = reinterpret_cast<const ulonglong2*>(a.uid_lo + i0);
            x = __ldcs(p); y = __ldcs(p + 1);
            lo_[0] = x.x; lo_[1] = x.y; lo_[2] = y.x; lo_[3] = y.y;
        }
        // keep bits of this thread's four events (bit k = event i0+k)
        u32 keep_bits = 0xFu;
        if (FILTERED) keep_bits = (a.keep[i0 >> 5] >> (i0 & 31)) & 0xFu;
        // first probe of all kept events in flight before anything else is looked at
        u32 sl[4];
        ulonglong2 pv[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            sl[k] = uid_home_slot(hi_[k], lo_[k], a.table_mask);
            pv[k] = make_ulonglong2(0, 0);
            if (!FILTERED || ((keep_bits >> k) & 1u)) pv[k] = ld_slot(a.table + sl[k], pol);
        }
        {
            const u32 pw = __ldcs(reinterpret_cast<const u32*>(a.phase + i0));
            const u32 tw = __ldcs(reinterpret_cast<const u32*>(a.type + i0));
#pragma unroll
            for (int k = 0; k < 4; k++) { ph[k] = (pw >> (8 * k)) & 0xFF; ty[k] = (tw >> (8 * k)) & 0xFF; }
        }
        if (!FILTERED) {
            // replay order check: rv must not decrease along arrival order (filter_flag_kernel does it for FILTERED)
            const ulonglong2* p = reinterpret_cast<const ulonglong2*>(a.rv + i0);
            const ulonglong2 x = __ldcs(p), y = __ldcs(p + 1);
            u64 prev = __shfl_up_sync(0xffffffffu, y.y, 1);
            if (lane == 0) prev = i0 > 0 ? a.rv[i0 - 1] : 0;
            if ((i0 < a.n && x.x < prev) || (i0 + 1 < a.n && x.y < x.x) || (i0 + 2 < a.n && y.x < x.y) ||
                (i0 + 3 < a.n && y.y < y.x))
                flags |= kFlagUnordered;
        }
        bool keep[4];
        u32 cnt = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            bool ok = i0 + k < a.n && ((keep_bits >> k) & 1u);
            if (ok && (ty[k] > kTypeDeleted || ph[k] > kPhMax)) { flags |= kFlagBadEvent; ok = FILTERED; }   // the run fails anyway
            if (ok && (hi_[k] | lo_[k]) == 0) { flags |= kFlagBadEvent; ok = FILTERED; }
            keep[k] = ok;
            cnt += ok;
        }
        u64 dst = i0;
        if (FILTERED) {
            // stable compaction behind the block's scanned offset
            const u32 block = tile / kSub;
            if (block != cur_block) { cur_block = block; run = a.block_off[block]; }
            u32 total;
            const u32 local = excl_scan_256(cnt, s_tot, &total);
            dst = (u64)run + local;
            run += total;
        }
        u64 rec[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            rec[k] = 0;
            if (keep[k]) {
                const u32 slot = intern_uid(a.table, a.table_mask, hi_[k], lo_[k], sl[k], pv[k], inserted, &a.ctr->flags, pol);
                rec[k] = rec_pack(a.L, slot, i0 + k, ty[k] & 3u, eff_phase(ph[k]) & 63u);   // the diff stage compares effective phases
                del_surv += ty[k] == kTypeDeleted;
#pragma unroll
                for (int p = 0; p < 4; p++)
                    if (p < a.key_passes) atomicAdd(&s_hist[p * kRadix + ((slot >> (8 * p)) & 0xFF)], 1u);
                if (FILTERED) __stcs(a.rec_out + dst++, rec[k]);
            }
        }
        if (!FILTERED) {
            // buffers are padded to a tile multiple, so the full 32 bytes are always writable
            ulonglong2* o = reinterpret_cast<ulonglong2*>(a.rec_out + i0);
            __stcs(o, make_ulonglong2(rec[0], rec[1]));       // streaming: leave L2 to the dictionary
            __stcs(o + 1, make_ulonglong2(rec[2], rec[3]));
        }
    }
    __syncthreads();
    for (int k = tid; k < a.key_passes * kRadix; k += kInternThreads)
        if (s_hist[k]) atomicAdd(&a.hist[k], s_hist[k]);
    // one atomic per warp for the run-wide counters
    inserted = __reduce_add_sync(0xffffffffu, inserted);
    del_surv = __reduce_add_sync(0xffffffffu, del_surv);
    flags = __reduce_or_sync(0xffffffffu, flags);
    if (lane == 0) {
        if (inserted) atomicAdd(&a.ctr->n_pods, inserted);
        if (del_surv) atomicAdd(&a.ctr->n_del_survived, (u64)del_surv);
        if (flags) atomicOr(&a.ctr->flags, flags);
    }
}

// The reference's two predicates for every event -> keep bitmap (one bit per arrival position), plus the
// replay-order check.  For the R-trace (dedupe off) this bitmap already is the emit bitmap: nothing has to
// be interned or sorted.  For filtered T-trace runs it drives the compaction in intern_pack_kernel.
// Four events per thread (16-byte loads); eight threads' nibbles make one bitmap word.
__global__ void __launch_bounds__(256) filter_flag_kernel(InternArgs a, u32* __restrict__ bitmap) {
    const u32 lane = threadIdx.x & 31;
    u32 flags = 0, surv = 0, del = 0;
    const u64 n4 = (a.n + 3) / 4;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    // warp-uniform loop bound: every lane takes part in the shuffles
    for (u64 wb = (u64)blockIdx.x * blockDim.x + (threadIdx.x & ~31u); wb < n4; wb += stride) {
        const u64 i0 = (wb + lane) * 4;
        const ulonglong2* p = reinterpret_cast<const ulonglong2*>(a.rv + i0);       // padded: readable past n
        const ulonglong2 x = __ldcs(p), y = __ldcs(p + 1);
        const uint4 q = __ldcs(reinterpret_cast<const uint4*>(a.ns + i0));
        const u32 pw = __ldcs(reinterpret_cast<const u32*>(a.phase + i0));
        const u32 tw = __ldcs(reinterpret_cast<const u32*>(a.type + i0));
        u64 prev = __shfl_up_sync(0xffffffffu, y.y, 1);
        if (lane == 0) prev = i0 > 0 ? a.rv[i0 - 1] : 0;
        if ((i0 < a.n && x.x < prev) || (i0 + 1 < a.n && x.y < x.x) || (i0 + 2 < a.n && y.x < x.y) ||
            (i0 + 3 < a.n && y.y < y.x))
            flags |= kFlagUnordered;
        const u32 ns[4] = {q.x, q.y, q.z, q.w};
        u32 nib = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const u32 ty = (tw >> (8 * k)) & 0xFF, ph = (pw >> (8 * k)) & 0xFF;
            if (i0 + k >= a.n) continue;
            if (ty > kTypeDeleted || ph > kPhMax) { flags |= kFlagBadEvent; continue; }
            if (passes_filters(a, ty, ph, ns[k])) { nib |= 1u << k; surv++; del += ty == kTypeDeleted; }
        }
        u32 w = nib << (4 * (lane & 7));
        w |= __shfl_xor_sync(0xffffffffu, w, 1);
        w |= __shfl_xor_sync(0xffffffffu, w, 2);
        w |= __shfl_xor_sync(0xffffffffu, w, 4);
        if ((lane & 7) == 0 && i0 < a.n) bitmap[i0 >> 5] = w;
    }
    surv = __reduce_add_sync(0xffffffffu, surv);
    del = __reduce_add_sync(0xffffffffu, del);
    flags = __reduce_or_sync(0xffffffffu, flags);
    if (lane == 0) {
        if (surv) atomicAdd(&a.ctr->n_survived, (u64)surv);
        if (del) atomicAdd(&a.ctr->n_tombstones, (u64)del);
        if (flags) atomicOr(&a.ctr->flags, flags);
    }
}

// ---- segmented adjacent-diff -> emit bitmap ----------------------------------
// Input: survivors sorted by (key, position).  A record is emitted iff it is a DELETED, opens its
// key's segment, follows a DELETED of the same key, or carries a different effective phase than
// its predecessor -- the bulk form of the sequential last-known-phase rule.
// Emission sets bit `position` of a bitmap (one bit per record of the arrival-ordered buffer;
// 1 B events = 125 MB, which the 126 MB L2 mostly keeps), so that the payload stage can walk
// arrival order and needs no second sort.  Two records per thread (one 16-byte load).
// One launch covers the positions [pos_lo, pos_hi): the host splits the position space into rounds so
// that the slice of the bitmap a round touches stays resident in L2 (re-reading the records for
// every round is cheaper than letting the atomics miss).
// The diff stage is bound by L2 atomics, one per flipped bit.  With uniform phase churn 80 % of the
// records are emitted; with a quiet cluster almost none are.  So the stage flips the MINORITY: a sample
// of the sorted records (whole key segments from the front of the array) estimates the emit fraction,
// and if it is above one half the bitmap starts as "every survivor emitted" and the diff clears bits.
// The estimate only picks the cheaper of two exact procedures; it cannot change the result.
__global__ void __launch_bounds__(256)
diff_sample_kernel(const u64* __restrict__ in, const u64* __restrict__ n_ptr, RecLayout L, Counters* __restrict__ ctr, u64 sample) {
    const u64 n = min(*n_ptr, sample);
    const u64 diff_mask = (~0ull << L.key_shift()) | 0x3Full;
    u32 emit = 0, total = 0;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        const u64 cur = in[i], p = i ? in[i - 1] : ~cur;
        emit += (((p ^ cur) & diff_mask) | ((p | cur) & 0x80ull)) != 0;
        total++;
    }
    emit = __reduce_add_sync(0xffffffffu, emit);
    total = __reduce_add_sync(0xffffffffu, total);
    if ((threadIdx.x & 31) == 0 && total) { atomicAdd(&ctr->sample_emit, emit); atomicAdd(&ctr->sample_total, total); }
}

// Start value of the emit bitmap: zeros, or (inverted) one bit per survivor -- the keep bitmap of a
// filtered run, else all positions below n.  Also latches the polarity for the diff launches.
__global__ void __launch_bounds__(256)
bitmap_init_kernel(u32* __restrict__ bitmap, const u32* __restrict__ keep, u64 n, Counters* __restrict__ ctr, int allow_invert) {
    const u32 inv = allow_invert && ctr->sample_total && 2ull * ctr->sample_emit > ctr->sample_total;
    if (blockIdx.x == 0 && threadIdx.x == 0) ctr->invert = inv;
    const u64 words = (n + 31) / 32;
    for (u64 w = (u64)blockIdx.x * blockDim.x + threadIdx.x; w < words; w += (u64)gridDim.x * blockDim.x) {
        u32 v = 0;
        if (inv) {
            v = keep ? keep[w] : 0xffffffffu;
            if (w == words - 1 && (n & 31)) v &= (1u << (n & 31)) - 1;
        }
        bitmap[w] = v;
    }
}

// CARRY: the dictionary and one state byte per slot (0 = no live phase known, else 1 + effective phase)
// survive from the previous batches of a stream.  The first record of a key then compares against that
// state instead of being emitted unconditionally, and the last record of a key leaves the new state.
// Heads read state_old, tails write state_new (a copy of state_old made before the launch), so the
// rounds and the threads of one round cannot see each other's updates.
// DEFER: the launch covers two slices of the position space with ONE read of the records.  Flips in
// [pos_lo, pos_hi) go straight to the bitmap; flips in [pos_hi, defer_hi) are appended to a list private to
// the warp (positions, 4 bytes each, in a record buffer the sort no longer needs; one warp scan per 128
// records, the cursor lives in a register) and diff_apply_kernel plays them back afterwards, when that slice
// of the bitmap can have the L2 to itself.  (Lists private to a thread need no scan, but their 4-byte appends
// turn into read-modify-write traffic in DRAM: measured 4.4 ms against 3.7 ms for two plain rounds.)
// Every DELETED survivor is emitted, so the tombstone count is the DELETED count intern_pack_kernel took.
struct DiffArgs {
    const u64* in; const u64* n_ptr;
    u32* bitmap; RecLayout L; Counters* ctr;
    u32 pos_lo, pos_hi, defer_hi;
    u32* defer_list; u32* defer_cnt; u64 region_cap;     // DEFER: region_cap entries per warp
    const u8* state_old; u8* state_new;                 // CARRY
};
#ifndef PTE_DIFF_PER_THREAD
#define PTE_DIFF_PER_THREAD 8      // measured on B200, diff stage of 1 B records: 2/4/8/12/16 -> 5.08/3.52/2.92/2.82/3.10 ms
#endif
static constexpr int kDiffPerThread = PTE_DIFF_PER_THREAD, kDiffThreads = 256;      // records per thread and iteration (even)

// A position is flipped at most once per run (one record per position), so XOR serves both polarities.
// Plain RED: keeps the compiler from wrapping the atomic in its warp-aggregation sequence (two flips of one
// warp almost never share a word here, and the sequence was a quarter of the kernel's instructions).
__device__ __forceinline__ void flip_bit(u32* __restrict__ bitmap, u32 pos) {
    asm volatile("red.global.xor.b32 [%0], %1;" ::"l"(bitmap + (pos >> 5)), "r"(1u << (pos & 31)) : "memory");
}

template <bool CARRY, bool DEFER>
__global__ void __launch_bounds__(kDiffThreads) diff_flag_kernel(DiffArgs a) {
    const u64 n = *a.n_ptr;
    const u32 lane = threadIdx.x & 31;
    // records carry the EFFECTIVE phase, and DELETED (type 2) is the only type with bit 7 set, so
    //   emit  <=>  key or phase differs from the predecessor, or either of the two is a DELETED
    const u32 ks = a.L.key_shift();
    const u64 diff_mask = (~0ull << ks) | 0x3Full;
    const u32 idx_mask = (u32)a.L.idx_mask();
    const u32 span = a.pos_hi - a.pos_lo, span_all = (DEFER ? a.defer_hi : a.pos_hi) - a.pos_lo;
    const bool invert = a.ctr->invert != 0;
    const u64 gwarp = ((u64)blockIdx.x * kDiffThreads + threadIdx.x) >> 5;
    u32* tail = DEFER ? a.defer_list + gwarp * a.region_cap : nullptr;     // end of this warp's list (same in all lanes)
    const u64 stride = (u64)gridDim.x * kDiffThreads * kDiffPerThread;
    // the loop bound is warp-uniform (wb = first record of the warp) so that every lane reaches the shuffles
    for (u64 wb = ((u64)blockIdx.x * kDiffThreads + (threadIdx.x & ~31u)) * kDiffPerThread; wb < n; wb += stride) {
        const u64 i = wb + lane * kDiffPerThread;
        const ulonglong2* src = reinterpret_cast<const ulonglong2*>(a.in + i);      // padded: readable past n
        u64 r[kDiffPerThread + 1];
#pragma unroll
        for (int k = 0; k < kDiffPerThread / 2; k++) {
            const ulonglong2 c = __ldcs(src + k);
            r[2 * k] = c.x; r[2 * k + 1] = c.y;
        }
        u64 p = __shfl_up_sync(0xffffffffu, r[kDiffPerThread - 1], 1);
        if (lane == 0) p = i > 0 ? a.in[i - 1] : ~r[0];                             // ~r[0]: a different key
        r[kDiffPerThread] = 0;
        if (CARRY) {
            r[kDiffPerThread] = __shfl_down_sync(0xffffffffu, r[0], 1);
            if (lane == 31) r[kDiffPerThread] = a.in[i + kDiffPerThread];
        }
        const u32 left = i >= n ? 0u : (n - i >= (u64)kDiffPerThread ? (u32)kDiffPerThread : (u32)(n - i));   // records of this lane
        u32 dpos[kDiffPerThread], dmask = 0;
#pragma unroll
        for (int k = 0; k < kDiffPerThread; k++) {
            const u64 cur = r[k];
            bool emit = (((p ^ cur) & diff_mask) | ((p | cur) & 0x80ull)) != 0;
            const u32 rel = ((u32)(cur >> 8) & idx_mask) - a.pos_lo;
            const bool mine = (u32)k < left && rel < span_all;
            if (CARRY && mine) {
                const u32 slot = (u32)(cur >> ks);
                if ((p >> ks) != (cur >> ks)) {                       // first record of this key in the batch
                    const u32 st = a.state_old[slot];
                    emit = (cur & 0x80ull) || st == 0 || st - 1 != (u32)(cur & 0x3F);
                }
                if (i + k + 1 >= n || (r[k + 1] >> ks) != (cur >> ks))   // last record of this key in the batch
                    a.state_new[slot] = (cur & 0x80ull) ? (u8)0 : (u8)(1 + (cur & 0x3F));
            }
            if (mine && emit != invert) {                  // flip the minority's bits only
                if (!DEFER || rel < span) flip_bit(a.bitmap, rel + a.pos_lo);
                else { dpos[k] = rel + a.pos_lo; dmask |= 1u << k; }
            }
            p = cur;
        }
        if (DEFER) {
            // append this warp's deferred positions to its list
            const u32 c = __popc(dmask);
            u32 incl = c;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const u32 t = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= (u32)d) incl += t;
            }
            u32* q = tail + (incl - c);
#pragma unroll
            for (int k = 0; k < kDiffPerThread; k++)
                if ((dmask >> k) & 1u) *q++ = dpos[k];
            tail += __shfl_sync(0xffffffffu, incl, 31);
        }
    }
    if (DEFER && lane == 0) a.defer_cnt[gwarp] = (u32)(tail - (a.defer_list + gwarp * a.region_cap));
}

// Plays the deferred flips back: list r holds cnt[r] positions.
__global__ void __launch_bounds__(256)
diff_apply_kernel(const u32* __restrict__ list, const u32* __restrict__ cnt, u64 region_cap, u32 n_regions, u32* __restrict__ bitmap) {
    for (u32 r = blockIdx.x; r < n_regions; r += gridDim.x) {
        const u32 c = cnt[r];
        const u32* l = list + (u64)r * region_cap;
        for (u32 j = threadIdx.x; j < c; j += blockDim.x) flip_bit(bitmap, __ldcs(l + j));
    }
}

// (type << 6 | phase) per event -> the two columns, 16 events per thread and step (the buffers are padded to a multiple of 16)
__global__ void __launch_bounds__(256) unpack_meta_kernel(u8* __restrict__ phase, u8* __restrict__ type, u64 n) {
    const u64 n16 = (n + 15) / 16;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (u64)gridDim.x * blockDim.x) {
        uint4 m = reinterpret_cast<const uint4*>(phase)[i];
        uint4 t;
        t.x = (m.x >> 6) & 0x03030303u; t.y = (m.y >> 6) & 0x03030303u; t.z = (m.z >> 6) & 0x03030303u; t.w = (m.w >> 6) & 0x03030303u;
        m.x &= 0x3F3F3F3Fu; m.y &= 0x3F3F3F3Fu; m.z &= 0x3F3F3F3Fu; m.w &= 0x3F3F3F3Fu;
        reinterpret_cast<uint4*>(phase)[i] = m;
        reinterpret_cast<uint4*>(type)[i] = t;
    }
}

// ---- a dictionary that lives as long as the watch (pod_watcher.py:264-269 never ends) ------------------------
// A stream's dictionary only ever gains uids: a DELETED pod keeps its slot with state 0.  When it fills up, the
// slots whose state byte is not 0 (pods with a live last-known phase) are re-inserted into a fresh table and their
// state moves with them; everything else is forgotten, which changes nothing: state 0 and "never seen" are the
// same to the transition rule.
__global__ void __launch_bounds__(256)
rebuild_table_kernel(const Slot* __restrict__ old_table, const u8* __restrict__ old_state, u64 cap, Slot* __restrict__ new_table,
                     u8* __restrict__ new_state, u32 mask, Counters* __restrict__ ctr) {
    const u64 pol = l2_policy_evict_last();
    u32 inserted = 0;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < cap; i += (u64)gridDim.x * blockDim.x) {
        const u32 st = old_state[i];
        if (st == 0) continue;
        const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(old_table + i);
        const u32 home = uid_home_slot(v.x, v.y, mask);
        const u32 sl = intern_uid(new_table, mask, v.x, v.y, home, ld_slot(new_table + home, pol), inserted, &ctr->flags, pol);
        new_state[sl] = (u8)st;
    }
    inserted = __reduce_add_sync(0xffffffffu, inserted);
    if ((threadIdx.x & 31) == 0 && inserted) atomicAdd(&ctr->n_pods, inserted);
}

// ---- payload columns --------------------------------------------------------
// Emitted records -> the columns a sink needs (idx, rv, key, ns_id, phase, type), in replay order.
struct AssembleArgs {
    u64 n;                 // positions = events in the log
    const u32* bitmap;     // emit bit per position
    const u64* rv; const u32* ns; const u8* phase; const u8* type;   // the log
    Counters* ctr;
    u32* lookback;
    u32* out_idx; u64* out_rv; u32* out_ns; u8* out_phase; u8* out_type;
};

// The interned key is not written here: few sinks want it, and pte_fetch can look it up on demand.

// Interned key of each emitted record, on demand: a read-only walk of the uid dictionary.
__global__ void __launch_bounds__(256)
key_lookup_kernel(const u32* __restrict__ out_idx, u64 n, const u64* __restrict__ uid_hi, const u64* __restrict__ uid_lo,
                  const Slot* __restrict__ table, u32 mask, u64* __restrict__ out_key) {
    for (u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (u64)gridDim.x * blockDim.x) {
        const u32 i = out_idx[j];
        const u64 hi = uid_hi[i], lo = uid_lo[i];
        const u32 home = uid_home_slot(hi, lo, mask);
        u64 key = ~0ull;
        for (int p = 0; p < kMaxProbe; ++p) {
            const u32 sl = probe_slot(home, p, mask);
            ulonglong2 v = *reinterpret_cast<const ulonglong2*>(table + sl);
            if (v.x == hi && v.y == lo) { key = sl; break; }
            if ((v.x | v.y) == 0) break;
        }
        out_key[j] = key;
    }
}

// Stable compaction of the positions whose emit bit is set (T-trace: set by diff_flag_kernel; R-trace: by
// filter_flag_kernel).  A tile is 256 bitmap words
// (8192 positions), one word per thread.  The bitmap is small (1 bit per event), so it is read twice:
// first to count every tile and scan the counts (two tiny kernels), then by the kernel below, whose
// tiles are independent of each other -- no look-back chain in the kernel that moves the bytes.

__global__ void __launch_bounds__(256)
bitmap_count_kernel(const u32* __restrict__ bitmap, u64 n, u32* __restrict__ tile_count) {
    __shared__ u32 s_w[8];
    const u32 n_tiles = (u32)((n + kAsmTile - 1) / kAsmTile);
    for (u32 tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const u64 w = (u64)tile * 256 + threadIdx.x;
        u32 c = w * 32 < n ? __popc(bitmap[w]) : 0u;
        c = __reduce_add_sync(0xffffffffu, c);
        if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = c;
        __syncthreads();
        if (threadIdx.x == 0) {
            u32 t = 0;
#pragma unroll
            for (int k = 0; k < 8; k++) t += s_w[k];
            tile_count[tile] = t;
        }
        __syncthreads();
    }
}

// exclusive scan of the tile counts, in place, by one block; also the total = n_emitted
__global__ void __launch_bounds__(1024) tile_scan_kernel(u32* __restrict__ tile_count, u64 n, Counters* __restrict__ ctr) {
    __shared__ u32 s_part[1024];
    const u32 n_tiles = (u32)((n + kAsmTile - 1) / kAsmTile);
    const u32 per = (n_tiles + 1023) / 1024;
    const u32 lo = min(n_tiles, threadIdx.x * per), hi = min(n_tiles, lo + per);
    u32 sum = 0;
    for (u32 t = lo; t < hi; t++) sum += tile_count[t];
    s_part[threadIdx.x] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        u32 run = 0;
        for (int k = 0; k < 1024; k++) { u32 v = s_part[k]; s_part[k] = run; run += v; }
        ctr->n_emitted = run;
    }
    __syncthreads();
    u32 run = s_part[threadIdx.x];
    for (u32 t = lo; t < hi; t++) { u32 v = tile_count[t]; tile_count[t] = run; run += v; }
}

#ifndef PTE_ASM_BLOCKS
#define PTE_ASM_BLOCKS 4         // measured: (rows, CTAs/SM) (4,6) 7.5 ms with spills, (4,4) 5.9 ms, (8,3) 5.8 ms
#endif
__global__ void __launch_bounds__(256, PTE_ASM_BLOCKS) assemble_bitmap_kernel(AssembleArgs a) {
    __shared__ unsigned short s_pos[kAsmTile];
    __shared__ u32 s_tot[8];
    const u32 tid = threadIdx.x;
    const u64 n = a.n;
    const u32 n_tiles = (u32)((n + kAsmTile - 1) / kAsmTile);
    for (u32 tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const u64 wbase = (u64)tile * 256 + tid;          // bitmap word of this thread
        u32 w = wbase * 32 < n ? __ldcs(a.bitmap + wbase) : 0u;
        u32 cnt;
        u32 local = excl_scan_256(__popc(w), s_tot, &cnt);
        const u32 p0 = tid * 32;
        while (w) {
            const u32 b = __ffs(w) - 1;
            w &= w - 1;
            s_pos[local++] = (unsigned short)(p0 + b);
        }
        __syncthreads();
        const u64 tbase = (u64)tile * kAsmTile, obase = a.lookback[tile];      // tile offsets from tile_scan_kernel
        // kAsmRows rows per thread and iteration, all gathers issued before the first store.  Only the arrival index
        // and the event type are written here (what a host that holds the log needs, pod_watcher.py:232-236); rv,
        // namespace and phase of the emitted records are gathered on demand (emitted_columns_kernel).
        for (u32 j0 = tid; j0 < cnt; j0 += kAsmRows * 256) {
            u32 idx[kAsmRows];
            u8 ty[kAsmRows];
#pragma unroll
            for (int k = 0; k < kAsmRows; k++) {
                const u32 j = j0 + k * 256;
                if (j < cnt) {
                    idx[k] = (u32)(tbase + s_pos[j]);
                    ty[k] = __ldcs(a.type + idx[k]);
                }
            }
#pragma unroll
            for (int k = 0; k < kAsmRows; k++) {
                const u32 j = j0 + k * 256;
                if (j < cnt) {
                    const u64 o = obase + j;
                    __stcs(a.out_idx + o, idx[k]);
                    a.out_type[o] = ty[k];
                }
            }
        }
        __syncthreads();      // s_pos is reused by the next tile
    }
}

// Order keys for a log that is declared to be in arrival order (pte_replay_pipelined with rv == NULL): 1, 2, 3, ...
__global__ void __launch_bounds__(256) iota_rv_kernel(u64* __restrict__ rv, u64 n, u64 first) {
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) rv[i] = first + i;
}

// rv, namespace and phase of the emitted records, on demand (pte_fetch of those columns, payload serialisation)
__global__ void __launch_bounds__(256)
emitted_columns_kernel(const u32* __restrict__ out_idx, u64 n, const u64* __restrict__ rv, const u32* __restrict__ ns,
                       const u8* __restrict__ phase, u64* __restrict__ out_rv, u32* __restrict__ out_ns, u8* __restrict__ out_phase) {
    for (u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (u64)gridDim.x * blockDim.x) {
        const u32 i = out_idx[j];
        out_rv[j] = rv[i]; out_ns[j] = ns[i]; out_phase[j] = phase[i];
    }
}

// ---- digest of the emitted records (parity tool, not part of the timed path) ----
__global__ void __launch_bounds__(256)
digest_kernel(const u32* __restrict__ out_idx, const u8* __restrict__ phase, const u8* __restrict__ out_type,
              const u64* __restrict__ uid_hi, const u64* __restrict__ uid_lo, const u64* __restrict__ rv,
              const u32* __restrict__ ns, Counters* __restrict__ ctr) {
    const u64 n = ctr->n_emitted;
    u64 acc = 0;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        u32 j = out_idx[i];
        acc += rec_digest(i, rv[j], uid_hi[j], uid_lo[j], ns[j], phase[j], out_type[i]);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0 && acc) atomicAdd(&ctr->digest, acc);
}

// ---- payload serialisation ------------------------------------------------------------------------
// The JSON body the reference would POST for an emitted record of a column log
// (json.dumps(_extract_pod_data(pod) + event_type), pod_watcher.py:159-202,:233; clusterapi_client.py:36).
// Everything that needs escaping is prepared by the host as ready-made JSON fragments; a record is
//   seg0 uid[:13] seg1 NS seg2 uid seg3 PHASE seg4 TYPE seg5
// with NS / PHASE / TYPE looked up by code.  One kernel sizes the records, the tile sums are scanned,
// and one warp per record then copies the fragments and formats the uid.
struct PayloadTable {
    const char* blob;      // all fragments, back to back
    const u32* off;        // fragment i = blob[off[i], off[i+1]):  6 segments, 3 types, n_phase phases, n_ns namespaces
    u32 n_phase, n_ns;
    __device__ __forceinline__ u32 len(u32 i) const { return off[i + 1] - off[i]; }
    __device__ __forceinline__ u32 type_frag(u32 t) const { return 6 + t; }
    __device__ __forceinline__ u32 phase_frag(u32 p) const { return 9 + min(p, n_phase - 1); }
    __device__ __forceinline__ u32 ns_frag(u32 n) const { return 9 + n_phase + min(n, n_ns - 1); }
};
static constexpr int kPayTile = 1024;      // records per scan tile

__device__ __forceinline__ u32 payload_len(const PayloadTable& t, u32 fixed, u32 ns, u32 ph, u32 ty) {
    return fixed + 13 + 36 + t.len(t.ns_frag(ns)) + t.len(t.phase_frag(ph)) + t.len(t.type_frag(ty));
}

__global__ void __launch_bounds__(256)
payload_len_kernel(PayloadTable t, const u32* __restrict__ ns, const u8* __restrict__ phase, const u8* __restrict__ type,
                   u64 first, u64 count, u32* __restrict__ tile_sum) {
    __shared__ u32 s_w[8];
    u32 fixed = 0;
    for (int k = 0; k < 6; k++) fixed += t.len(k);
    const u64 n_tiles = (count + kPayTile - 1) / kPayTile;
    for (u64 tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        u32 sum = 0;
        for (u32 k = threadIdx.x; k < kPayTile; k += 256) {
            const u64 j = tile * kPayTile + k;
            if (j < count) sum += payload_len(t, fixed, ns[first + j], phase[first + j], type[first + j]);
        }
        sum = __reduce_add_sync(0xffffffffu, sum);
        if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = sum;
        __syncthreads();
        if (threadIdx.x == 0) {
            u32 v = 0;
            for (int k = 0; k < 8; k++) v += s_w[k];
            tile_sum[tile] = v;
        }
        __syncthreads();
    }
}

// exclusive scan of n 32-bit tile sums by one block -> 64-bit offsets (n + 1 entries)
__global__ void __launch_bounds__(1024) scan_u32_to_u64_kernel(const u32* __restrict__ in, u64 n, u64* __restrict__ out) {
    __shared__ u64 s_part[1024];
    const u64 per = (n + 1023) / 1024;
    const u64 lo = min(n, threadIdx.x * per), hi = min(n, lo + per);
    u64 sum = 0;
    for (u64 i = lo; i < hi; i++) sum += in[i];
    s_part[threadIdx.x] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        u64 run = 0;
        for (int k = 0; k < 1024; k++) { u64 v = s_part[k]; s_part[k] = run; run += v; }
        out[n] = run;
    }
    __syncthreads();
    u64 run = s_part[threadIdx.x];
    for (u64 i = lo; i < hi; i++) { out[i] = run; run += in[i]; }
}

__device__ __forceinline__ void warp_copy(char* __restrict__ dst, const char* __restrict__ src, u32 len, u32 lane) {
    for (u32 k = lane; k < len; k += 32) dst[k] = src[k];
}
__device__ __forceinline__ char uid_char(u64 hi, u64 lo, u32 pos) {      // pos in the 8-4-4-4-12 text form
    if (pos == 8 || pos == 13 || pos == 18 || pos == 23) return '-';
    const u32 nib = pos - (pos > 8) - (pos > 13) - (pos > 18) - (pos > 23);      // 0..31, most significant first
    const u64 w = nib < 16 ? hi : lo;
    const u32 v = (u32)(w >> (60 - 4 * (nib & 15))) & 15u;
    return (char)(v < 10 ? '0' + v : 'a' + v - 10);
}

__global__ void __launch_bounds__(256)
payload_write_kernel(PayloadTable t, const u32* __restrict__ out_idx, const u32* __restrict__ ns, const u8* __restrict__ phase,
                     const u8* __restrict__ type, const u64* __restrict__ uid_hi, const u64* __restrict__ uid_lo,
                     u64 first, u64 count, const u64* __restrict__ tile_off, char* __restrict__ out, u64* __restrict__ rec_off) {
    __shared__ u32 s_len[kPayTile];
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    u32 fixed = 0;
    for (int k = 0; k < 6; k++) fixed += t.len(k);
    const u64 n_tiles = (count + kPayTile - 1) / kPayTile;
    for (u64 tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        for (u32 k = threadIdx.x; k < kPayTile; k += 256) {
            const u64 j = tile * kPayTile + k;
            s_len[k] = j < count ? payload_len(t, fixed, ns[first + j], phase[first + j], type[first + j]) : 0u;
        }
        __syncthreads();
        // each warp owns 128 consecutive records of the tile: running offset inside the warp's slice
        u64 base = tile_off[tile];
        for (u32 k = 0; k < warp * 128; k += 32) base += __reduce_add_sync(0xffffffffu, s_len[k + lane]);
        for (u32 r = 0; r < 128; r++) {
            const u32 k = warp * 128 + r;
            const u64 j = tile * kPayTile + k;
            if (j >= count) break;
            const u64 e = first + j;
            if (lane == 0) rec_off[j] = base;
            char* d = out + base;
            const u32 i = out_idx[e];
            const u64 hi = uid_hi[i], lo = uid_lo[i];
            u32 o = 0, f;
            warp_copy(d + o, t.blob + t.off[0], t.len(0), lane); o += t.len(0);
            if (lane < 13) d[o + lane] = uid_char(hi, lo, lane);
            o += 13;
            warp_copy(d + o, t.blob + t.off[1], t.len(1), lane); o += t.len(1);
            f = t.ns_frag(ns[e]); warp_copy(d + o, t.blob + t.off[f], t.len(f), lane); o += t.len(f);
            warp_copy(d + o, t.blob + t.off[2], t.len(2), lane); o += t.len(2);
            d[o + lane] = uid_char(hi, lo, lane);
            if (lane < 4) d[o + 32 + lane] = uid_char(hi, lo, 32 + lane);
            o += 36;
            warp_copy(d + o, t.blob + t.off[3], t.len(3), lane); o += t.len(3);
            f = t.phase_frag(phase[e]); warp_copy(d + o, t.blob + t.off[f], t.len(f), lane); o += t.len(f);
            warp_copy(d + o, t.blob + t.off[4], t.len(4), lane); o += t.len(4);
            f = t.type_frag(type[e]); warp_copy(d + o, t.blob + t.off[f], t.len(f), lane); o += t.len(f);
            warp_copy(d + o, t.blob + t.off[5], t.len(5), lane);
            base += s_len[k];
        }
        __syncthreads();
    }
}

// ---- logs that are not in resourceVersion order -------------------------------------------------
// Replay order is ascending rv (ties: arrival).  Such a log is brought into that order once: records
// (rv - rv_min, arrival index) go through the same radix sort, the columns are gathered into shadow
// buffers in sorted order, the normal pipeline runs on those, and the emitted indices are mapped back.
__global__ void __launch_bounds__(256) rv_minmax_kernel(const u64* __restrict__ rv, u64 n, u64* __restrict__ mm) {
    u64 lo = ~0ull, hi = 0;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        const u64 v = rv[i];
        lo = min(lo, v); hi = max(hi, v);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if ((threadIdx.x & 31) == 0) { atomicMin(mm, lo); atomicMax(mm + 1, hi); }
}

__global__ void __launch_bounds__(256)
rv_pack_kernel(const u64* __restrict__ rv, u64 n, u64 rv_min, u32 idx_bits, u64* __restrict__ rec) {
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x)
        rec[i] = ((rv[i] - rv_min) << idx_bits) | i;
}

struct LogColumns { u64 *uid_hi, *uid_lo, *rv; u32* ns; u8 *phase, *type; };

__global__ void __launch_bounds__(256)
permute_log_kernel(const u64* __restrict__ sorted, u64 n, u64 idx_mask, LogColumns src, LogColumns dst, u32* __restrict__ perm) {
    for (u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (u64)gridDim.x * blockDim.x) {
        const u64 i = sorted[j] & idx_mask;
        perm[j] = (u32)i;
        dst.uid_hi[j] = src.uid_hi[i]; dst.uid_lo[j] = src.uid_lo[i]; dst.rv[j] = src.rv[i];
        dst.ns[j] = src.ns[i]; dst.phase[j] = src.phase[i]; dst.type[j] = src.type[i];
    }
}

__global__ void __launch_bounds__(256)
map_idx_kernel(u32* __restrict__ out_idx, const Counters* __restrict__ ctr, const u32* __restrict__ perm) {
    const u64 n = ctr->n_emitted;
    for (u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (u64)gridDim.x * blockDim.x)
        out_idx[j] = perm[out_idx[j]];
}

// ---- synthetic replay log -----------------------------------------------------
struct GenArgs {
    GenParams g;
    u64* uid_hi; u64* uid_lo; u64* rv; u32* ns; u8* phase; u8* type;
    u64* rec_out;      // (q, j, draw) records, sorted by q afterwards to derive `type`
    RecLayout L;
};

__global__ void __launch_bounds__(256) gen_fill_kernel(GenArgs a) {
    const u64 n = a.g.n_events;
    for (u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (u64)gridDim.x * blockDim.x) {
        GenEvent e = gen_event(a.g, j);
        a.uid_hi[j] = e.uid_hi; a.uid_lo[j] = e.uid_lo; a.rv[j] = e.rv;
        a.ns[j] = e.ns; a.phase[j] = (u8)e.phase;
        a.rec_out[j] = (e.q << a.L.key_shift()) | (j << 8) | e.draw;
    }
}

__global__ void __launch_bounds__(256)
gen_type_kernel(const u64* __restrict__ rec, u64 n, RecLayout L, u8* __restrict__ type) {
    const u32 ks = L.key_shift();
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        u64 c = rec[i];
        bool has_prev = false;
        u32 prev_draw = 0;
        if (i > 0) {
            u64 p = rec[i - 1];
            has_prev = (p >> ks) == (c >> ks);
            prev_draw = (u32)(p & 1);
        }
        type[rec_idx(L, c)] = (u8)gen_type(has_prev, prev_draw, (u32)(c & 1));
    }
}
