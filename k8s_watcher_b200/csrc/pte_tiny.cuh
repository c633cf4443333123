// Small batches in ONE launch.
//
// The live watch loop (watcher/pod_watcher.py:264-269) hands the engine a few thousand events at a time; through the
// bulk kernels such a batch is ten launches, three memsets and a copy of the state bytes -- 114 us on the device and
// 233 us of wall time for 4 096 events, nearly all of it launch latency (profiles/r02_live_loop.txt).  For batches of
// up to kTinyMax events one CTA does the whole replay: filter, intern the uids in the (global) dictionary, sort the
// records by (pod, arrival) in shared memory, compare neighbours and the carried state, leave the new state, and
// compact the notifications in arrival order.  Semantics are those of the bulk path, kernel by kernel:
//   filter_flag_kernel / intern_pack_kernel<true>   -> phase 1
//   radix passes on the key                         -> phase 2 (two stable counting passes in shared memory on a 13-bit id of the slot;
//                                                      a bitonic network on the 64-bit records took 3/4 of the kernel's 57 us)
//   diff_flag_kernel<CARRY>                         -> phase 3 (state read for heads, barrier, state written by tails: in place)
//   bitmap_count / tile_scan / assemble_bitmap      -> phase 4
// A batch that fails (dictionary full, rv out of order, bad event code) leaves the state bytes untouched.
#pragma once
#include "pte_kernels.cuh"

static constexpr u32 kTinyMax = 4096, kTinyThreads = 1024, kTinyPer = kTinyMax / kTinyThreads;

struct TinyArgs {
    // the batch as the caller handed it over (src), and the engine's own columns (dst; may be the same memory)
    const u64 *src_hi, *src_lo, *src_rv; const u32* src_ns; const u8 *src_phase, *src_type;
    u64 *uid_hi, *uid_lo, *rv; u32* ns; u8 *phase, *type;
    u32 n;
    Slot* table; u32 table_mask;
    const u32* ns_bitmap; u32 ns_words; int ns_filter_on; int critical_only;
    int carry;
    u8* state;                 // carry: one byte per slot, updated in place
    u64 prev_rv;               // carry: last resourceVersion of the previous batch (0 for the first)
    Counters* ctr;
    u32* bitmap;               // emit bitmap of the batch, arrival order
    u32* out_idx; u8* out_type;
    u8* packed;                // Counters | out_idx[n_emitted] | out_type[n_emitted]: one device-to-host copy brings all of it
};

__device__ __forceinline__ u32 tiny_block_excl_scan(u32 v, u32* s_warp, u32* total) {
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    u32 inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const u32 t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= (u32)o) inc += t; }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        u32 w = s_warp[lane], winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const u32 t = __shfl_up_sync(0xffffffffu, winc, o); if (lane >= (u32)o) winc += t; }
        s_warp[lane] = winc - w;
        if (lane == 31) *total = winc;
    }
    __syncthreads();
    return s_warp[warp] + inc - v;
}

// Shared memory of the kernel (dynamic, 84 KB):
//   ids   [8192] u32   hash set slot -> small id (its position in the set); later the per-warp digit counts of the sort
//   keyA/B[4096] u32   (id << 12 | arrival index), ping-pong of the two radix passes
//   slot  [4096] u32   dictionary slot of event i        meta[4096] u8   type << 6 | effective phase of event i
static constexpr u32 kTinyIds = 8192;
static constexpr size_t kTinySmem = kTinyIds * 4 + 2 * kTinyMax * 4 + kTinyMax * 4 + kTinyMax;

// One stable counting pass over `bits` bits at `shift` of the keys (warp w owns keys 128 w .. 128 w + 127, lane l the
// keys 128 w + 32 k + l: round k of a warp comes before round k + 1, lanes in order inside a round).
__device__ __forceinline__ void tiny_radix_pass(const u32* __restrict__ in, u32* __restrict__ out, u32* __restrict__ cnt,
                                                u32* s_warp, u32* s_total, u32 shift, u32 bits) {
    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const u32 bins = 1u << bits, mask = bins - 1;
    for (u32 c = tid; c < bins * 32; c += kTinyThreads) cnt[c] = 0;          // cnt[bin * 32 + warp]
    __syncthreads();
    u32 key[kTinyPer], rank[kTinyPer];
#pragma unroll
    for (u32 k = 0; k < kTinyPer; k++) {
        key[k] = in[warp * 128 + k * 32 + lane];
        const u32 d = (key[k] >> shift) & mask;
        // (MATCH.ANY: eight per thread.  Peer masks through shared-memory atomics, as in the bulk radix pass, were measured
        //  slower here: 38 against 32 us for the kernel -- more barriers, twice the counters to clear)
        const u32 peers = __match_any_sync(0xffffffffu, d);
        const u32 leader = __ffs(peers) - 1;
        u32 before = 0;
        if (lane == leader) { before = cnt[d * 32 + warp]; cnt[d * 32 + warp] = before + __popc(peers); }
        before = __shfl_sync(0xffffffffu, before, leader);
        rank[k] = before + __popc(peers & ((1u << lane) - 1u));
        __syncwarp();
    }
    __syncthreads();
    // exclusive scan of the counts in (bin, warp) order: bins * 32 <= 4096 counters, up to four per thread
    u32 c4[4], sum = 0;
#pragma unroll
    for (u32 k = 0; k < 4; k++) { const u32 c = tid * 4 + k; c4[k] = c < bins * 32 ? cnt[c] : 0u; sum += c4[k]; }
    u32 off = tiny_block_excl_scan(sum, s_warp, s_total);
#pragma unroll
    for (u32 k = 0; k < 4; k++) { const u32 c = tid * 4 + k; if (c < bins * 32) cnt[c] = off; off += c4[k]; }
    __syncthreads();
#pragma unroll
    for (u32 k = 0; k < kTinyPer; k++) out[cnt[((key[k] >> shift) & mask) * 32 + warp] + rank[k]] = key[k];
    __syncthreads();
}

__global__ void __launch_bounds__(kTinyThreads) tiny_replay_kernel(TinyArgs a) {
    extern __shared__ __align__(16) u8 s_mem[];
    u32* s_ids = reinterpret_cast<u32*>(s_mem);
    u32* s_keyA = s_ids + kTinyIds;
    u32* s_keyB = s_keyA + kTinyMax;
    u32* s_slot = s_keyB + kTinyMax;
    u8* s_meta = reinterpret_cast<u8*>(s_slot + kTinyMax);
    __shared__ u32 s_bits[kTinyMax / 32];
    __shared__ u32 s_warp[32];
    __shared__ u32 s_flags, s_inserted, s_del, s_surv, s_total;
    const u32 tid = threadIdx.x;
    const u32 kNoKey = 0xFFFFFFFFu;              // id 2^20 - 1: sorts behind every real key (ids are below 8192)
    if (tid < kTinyMax / 32) s_bits[tid] = 0;
    if (tid == 0) { s_flags = 0; s_inserted = 0; s_del = 0; s_surv = 0; }
    for (u32 c = tid; c < kTinyIds; c += kTinyThreads) s_ids[c] = 0;
    __syncthreads();

    // ---- 1. columns in (and on into the engine's own arrays), predicates, interning, small ids ----
    const u64 pol = l2_policy_evict_last();
    u32 flags = 0, inserted = 0, del = 0, surv = 0;
    {
        // everything a thread needs of its (up to) four events is requested before anything is looked at: the loads of all
        // columns, then the first probe of every kept event -- one memory latency each instead of four in a row
        u64 hi[kTinyPer], lo[kTinyPer], rv[kTinyPer], before[kTinyPer];
        u32 ns[kTinyPer], ph[kTinyPer], ty[kTinyPer], home[kTinyPer];
        ulonglong2 pv[kTinyPer];
        bool keep[kTinyPer];
#pragma unroll
        for (u32 k = 0; k < kTinyPer; k++) {
            const u32 i = tid + k * kTinyThreads;
            const bool in = i < a.n;
            hi[k] = in ? a.src_hi[i] : 0; lo[k] = in ? a.src_lo[i] : 0; rv[k] = in ? a.src_rv[i] : 0;
            before[k] = in ? (i ? a.src_rv[i - 1] : a.prev_rv) : 0;
            ns[k] = in && a.src_ns ? a.src_ns[i] : 0u; ph[k] = in ? a.src_phase[i] : 0u; ty[k] = in ? a.src_type[i] : 0u;
        }
#pragma unroll
        for (u32 k = 0; k < kTinyPer; k++) {
            const u32 i = tid + k * kTinyThreads;
            keep[k] = false;
            pv[k] = make_ulonglong2(0, 0);
            home[k] = 0;
            if (i < a.n) {
                if (a.src_hi != a.uid_hi) { a.uid_hi[i] = hi[k]; a.uid_lo[i] = lo[k]; a.rv[i] = rv[k]; a.ns[i] = ns[k]; a.phase[i] = (u8)ph[k]; a.type[i] = (u8)ty[k]; }
                if (rv[k] < before[k]) flags |= kFlagUnordered;
                if (ty[k] > kTypeDeleted || ph[k] > kPhMax) flags |= kFlagBadEvent;       // (as filter_flag_kernel: whatever the filters say)
                else keep[k] = passes_filters(a, ty[k], ph[k], ns[k]);
                if (keep[k] && (hi[k] | lo[k]) == 0) { flags |= kFlagBadEvent; keep[k] = false; }
                if (keep[k]) { home[k] = uid_home_slot(hi[k], lo[k], a.table_mask); pv[k] = ld_slot(a.table + home[k], pol); }
            }
        }
#pragma unroll
        for (u32 k = 0; k < kTinyPer; k++) {
            const u32 i = tid + k * kTinyThreads;
            u32 key = kNoKey;
            if (keep[k]) {
                const u32 slot = intern_uid(a.table, a.table_mask, hi[k], lo[k], home[k], pv[k], inserted, &s_flags, pol);
                // the slot's place in a set of at most 4 096 slots is its id for the sort (13 bits instead of 25)
                u32 id = (slot * 0x9E3779B1u) >> 19;
                for (;;) {
                    const u32 old = atomicCAS(&s_ids[id], 0u, slot + 1);
                    if (old == 0 || old == slot + 1) break;
                    id = (id + 1) & (kTinyIds - 1);
                }
                key = (id << 12) | i;
                s_slot[i] = slot;
                s_meta[i] = (u8)(((ty[k] & 3u) << 6) | (eff_phase(ph[k]) & 63u));
                del += ty[k] == kTypeDeleted;
                surv++;
            }
            s_keyA[i] = key;
        }
    }
    inserted = __reduce_add_sync(0xffffffffu, inserted);
    del = __reduce_add_sync(0xffffffffu, del);
    surv = __reduce_add_sync(0xffffffffu, surv);
    flags = __reduce_or_sync(0xffffffffu, flags);
    if ((tid & 31) == 0) {
        if (inserted) atomicAdd(&s_inserted, inserted);
        if (del) atomicAdd(&s_del, del);
        if (surv) atomicAdd(&s_surv, surv);
        if (flags) atomicOr(&s_flags, flags);
    }
    __syncthreads();
    const bool failed = s_flags != 0;

    // ---- 2. stable sort on the id: two counting passes (7 + 7 bits; the 14th bit is set only in kNoKey) ----
    // The keys start in arrival order and both passes are stable, so equal ids end up in arrival order.
    if (!failed) {
        tiny_radix_pass(s_keyA, s_keyB, s_ids, s_warp, &s_total, 12, 7);
        tiny_radix_pass(s_keyB, s_keyA, s_ids, s_warp, &s_total, 19, 7);
    }

    // ---- 3. neighbours and the carried state -> emit bits; tails leave the new state ----
    u32 tail_slot[kTinyPer];
    u8 tail_state[kTinyPer];
    u32 tails = 0;
    if (!failed) {
        u32 st_[kTinyPer];
#pragma unroll
        for (u32 k = 0; k < kTinyPer; k++) {                               // the state bytes of the heads, all in flight together
            const u32 i = tid + k * kTinyThreads;
            const u32 cur = s_keyA[i], prev = i ? s_keyA[i - 1] : kNoKey;
            st_[k] = (a.carry && cur != kNoKey && (prev >> 12) != (cur >> 12)) ? a.state[s_slot[cur & 0xFFFu]] : 0u;
        }
#pragma unroll
        for (u32 k = 0; k < kTinyPer; k++) {
            const u32 i = tid + k * kTinyThreads;
            const u32 cur = s_keyA[i];
            if (cur == kNoKey) continue;
            const u32 prev = i ? s_keyA[i - 1] : kNoKey, next = i + 1 < kTinyMax ? s_keyA[i + 1] : kNoKey;
            const u32 pos = cur & 0xFFFu, slot = s_slot[pos], m = s_meta[pos];
            bool emit;
            if ((prev >> 12) != (cur >> 12)) {                            // first record of this pod in the batch
                const u32 st = st_[k];
                emit = (m & 0x80u) || st == 0 || st - 1 != (m & 0x3Fu);
            } else {
                const u32 pm = s_meta[prev & 0xFFFu];
                emit = (((pm ^ m) & 0x3Fu) | ((pm | m) & 0x80u)) != 0;
            }
            if (emit) atomicOr(&s_bits[pos >> 5], 1u << (pos & 31));
            if (a.carry && (next >> 12) != (cur >> 12)) {                 // last record of this pod in the batch
                tail_slot[k] = slot; tail_state[k] = (m & 0x80u) ? (u8)0 : (u8)(1 + (m & 0x3Fu));
                tails |= 1u << k;
            }
        }
    }
    __syncthreads();            // every head has read its state byte
#pragma unroll
    for (u32 k = 0; k < kTinyPer; k++)
        if ((tails >> k) & 1u) a.state[tail_slot[k]] = tail_state[k];

    // ---- 4. notifications in arrival order ----
    const u32 nib = (s_bits[tid >> 3] >> ((tid & 7) * 4)) & 0xFu;       // events 4 tid .. 4 tid + 3
    const u32 off = tiny_block_excl_scan(__popc(nib), s_warp, &s_total);
    u32* p_idx = reinterpret_cast<u32*>(a.packed + sizeof(Counters));
    u8* p_type = a.packed + sizeof(Counters) + (size_t)s_total * 4;
    u32 o = off;
#pragma unroll
    for (u32 k = 0; k < 4; k++) {
        if ((nib >> k) & 1u) {
            const u32 i = tid * 4 + k;
            const u8 ty = a.src_type[i];
            a.out_idx[o] = i; a.out_type[o] = ty;
            p_idx[o] = i; p_type[o] = ty;
            o++;
        }
    }
    if (tid < (a.n + 31) / 32) a.bitmap[tid] = s_bits[tid];
    if (tid == 0) {
        Counters c;
        memset(&c, 0, sizeof(c));
        c.n_survived = s_surv; c.n_emitted = s_total; c.n_del_survived = s_del;
        c.n_pods = s_inserted; c.flags = s_flags;
        *a.ctr = c;
        *reinterpret_cast<Counters*>(a.packed) = c;
    }
}
