// Device kernels of the replay path other than the sort:
//   intern_pack_kernel   filters (pod_watcher.py:204-212, :226-229) + uid interning + record packing
//   diff_compact_kernel  segmented adjacent-diff + ballot/popc compaction
//   assemble_kernel      payload columns of the emitted records, in replay order (pod_watcher.py:232-236)
//   digest_kernel        order-sensitive digest for large-scale parity
//   gen_* kernels        synthetic replay logs (spec in include/pte.h)
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
    u32 ticket[4];         // tile tickets: [0] intern, [1] radix pass (reset per pass), [2..3] spare
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

// uid -> slot of an open-addressing table (linear probing, 16-byte slots, 0:0 = empty).
// Exact: two uids share a slot only if all 128 bits are equal.
// A plain 16-byte load may see a slot as empty when it is not (stale line), or half-written
// (0:0 -> hi:lo is the only transition a slot ever makes).  Either shows up as a zero half,
// and then the 128-bit CAS -- which returns the true content -- decides.
__device__ __forceinline__ u32 intern_uid(Slot* __restrict__ table, u32 mask, u64 hi, u64 lo, u32& inserted,
                                          u32* __restrict__ flags) {
    u32 sl = (u32)mix64(hi ^ (lo * 0x9E3779B97F4A7C15ull)) & mask;
    for (int p = 0; p < kMaxProbe; ++p) {
        ulonglong2 v = *reinterpret_cast<const ulonglong2*>(table + sl);
        if (v.x == 0 || v.y == 0) {
            ulonglong2 old = cas128(table + sl, 0, 0, hi, lo);
            if (old.x == 0 && old.y == 0) { inserted++; return sl; }
            if (old.x == hi && old.y == lo) return sl;
        } else if (v.x == hi && v.y == lo) {
            return sl;
        }
        sl = (sl + 1) & mask;
    }
    atomicOr(flags, kFlagTableFull);
    return 0;
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
    u32* lookback;     // one word per tile (FILTERED only)
};

static constexpr int kInternThreads = 256;
static constexpr int kInternPerThread = 4;
static constexpr int kInternTile = kInternThreads * kInternPerThread;

// One event per (thread, k): thread t of a tile owns events 4t..4t+3, loaded with 16-byte vectors.
// FILTERED == false: no predicate is active, record i goes to rec_out[i].
// FILTERED == true : survivors are compacted in arrival order (tile totals chained by look-back).
template <bool FILTERED>
__global__ void __launch_bounds__(kInternThreads) intern_pack_kernel(InternArgs a) {
    __shared__ u32 s_tot[8];
    __shared__ u32 s_tile, s_tile_excl;
    const u32 tid = threadIdx.x, lane = tid & 31;
    u32 inserted = 0, del_surv = 0, flags = 0;
    if (!FILTERED && blockIdx.x == 0 && tid == 0) a.ctr->n_survived = a.n;

    for (;;) {
        __syncthreads();
        if (tid == 0) s_tile = atomicAdd(&a.ctr->ticket[0], 1u);
        __syncthreads();
        const u32 tile = s_tile;
        const u64 i0 = (u64)tile * kInternTile + (u64)tid * kInternPerThread;
        if ((u64)tile * kInternTile >= a.n) break;

        u64 hi[4], lo[4], rv[4];
        u32 ns[4], ph[4], ty[4];
        {
            const ulonglong2* p = reinterpret_cast<const ulonglong2*>(a.uid_hi + i0);
            ulonglong2 x = __ldcs(p), y = __ldcs(p + 1);
            hi[0] = x.x; hi[1] = x.y; hi[2] = y.x; hi[3] = y.y;
            p = reinterpret_cast<const ulonglong2*>(a.uid_lo + i0);
            x = __ldcs(p); y = __ldcs(p + 1);
            lo[0] = x.x; lo[1] = x.y; lo[2] = y.x; lo[3] = y.y;
            p = reinterpret_cast<const ulonglong2*>(a.rv + i0);
            x = __ldcs(p); y = __ldcs(p + 1);
            rv[0] = x.x; rv[1] = x.y; rv[2] = y.x; rv[3] = y.y;
            uint4 q = __ldcs(reinterpret_cast<const uint4*>(a.ns + i0));
            ns[0] = q.x; ns[1] = q.y; ns[2] = q.z; ns[3] = q.w;
            u32 pw = __ldcs(reinterpret_cast<const u32*>(a.phase + i0));
            u32 tw = __ldcs(reinterpret_cast<const u32*>(a.type + i0));
#pragma unroll
            for (int k = 0; k < 4; k++) { ph[k] = (pw >> (8 * k)) & 0xFF; ty[k] = (tw >> (8 * k)) & 0xFF; }
        }
        // replay order check: rv must not decrease along arrival order
        {
            u64 prev = __shfl_up_sync(0xffffffffu, rv[3], 1);
            if (lane == 0) prev = i0 > 0 ? a.rv[i0 - 1] : 0;
            if (i0 < a.n && rv[0] < prev) flags |= kFlagUnordered;
#pragma unroll
            for (int k = 1; k < 4; k++)
                if (i0 + k < a.n && rv[k] < rv[k - 1]) flags |= kFlagUnordered;
        }
        bool keep[4];
        u32 cnt = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            bool ok = i0 + k < a.n;
            if (ok && (ty[k] > kTypeDeleted || ph[k] > kPhMax || (hi[k] | lo[k]) == 0)) { flags |= kFlagBadEvent; ok = false; }
            if (FILTERED && ok) {
                // should_process_event, pod_watcher.py:209
                if (a.critical_only && ty[k] != kTypeDeleted && ph[k] != kPhNoStatus && ph[k] != kPhFailed &&
                    ph[k] != kPhSucceeded)
                    ok = false;
                // namespace allow-list, pod_watcher.py:227
                if (ok && a.ns_filter_on) {
                    u32 w = ns[k] >> 5;
                    ok = w < a.ns_words && ((__ldg(a.ns_bitmap + w) >> (ns[k] & 31)) & 1u);
                }
            }
            keep[k] = ok;
            cnt += ok;
        }
        u64 dst = i0;
        if (FILTERED) {
            u32 total;
            u32 local = excl_scan_256(cnt, s_tot, &total);
            if (tid == 0) {
                u32 excl = 0;
                if (tile == 0) {
                    st_relaxed_u32(a.lookback, kLbInclusive | total);
                } else {
                    st_relaxed_u32(a.lookback + tile, kLbPartial | total);
                    u32 t = tile - 1;
                    for (;;) {
                        u32 v = ld_relaxed_u32(a.lookback + t);
                        if ((v >> 30) == 0) { __nanosleep(20); continue; }
                        excl += v & kLbValue;
                        if ((v >> 30) == 2) break;
                        t--;
                    }
                    st_relaxed_u32(a.lookback + tile, kLbInclusive | (excl + total));
                }
                s_tile_excl = excl;
                if ((u64)(tile + 1) * kInternTile >= a.n) a.ctr->n_survived = excl + total;   // last tile
            }
            __syncthreads();
            dst = (u64)s_tile_excl + local;
        }
        u64 rec[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            rec[k] = 0;
            if (keep[k]) {
                u32 sl = intern_uid(a.table, a.table_mask, hi[k], lo[k], inserted, &a.ctr->flags);
                rec[k] = rec_pack(a.L, sl, i0 + k, ty[k], ph[k]);
                del_surv += ty[k] == kTypeDeleted;
            }
        }
        if (FILTERED) {
#pragma unroll
            for (int k = 0; k < 4; k++)
                if (keep[k]) a.rec_out[dst++] = rec[k];
        } else {
            // buffers are padded to a tile multiple, so the full 32 bytes are always writable
            ulonglong2* o = reinterpret_cast<ulonglong2*>(a.rec_out + i0);
            o[0] = make_ulonglong2(rec[0], rec[1]);
            o[1] = make_ulonglong2(rec[2], rec[3]);
        }
    }
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

// ---- segmented adjacent-diff + compaction -----------------------------------
// Input: survivors sorted by (key, arrival).  A record is emitted iff it is a DELETED, opens its
// key's segment, follows a DELETED of the same key, or carries a different effective phase than
// its predecessor -- the bulk form of the sequential last-known-phase rule.
static constexpr int kDiffThreads = 256;
static constexpr int kDiffItems = 8;

__global__ void __launch_bounds__(kDiffThreads)
diff_compact_kernel(const u64* __restrict__ in, u64* __restrict__ out, RecLayout L, Counters* __restrict__ ctr) {
    __shared__ u32 s_cnt[kDiffItems][kDiffThreads / 32];
    __shared__ u32 s_base;
    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const u32 lt = lanemask_lt();
    const u64 n = ctr->n_survived;
    const u32 ks = L.key_shift();
    constexpr int TILE = kDiffThreads * kDiffItems;
    u32 tomb = 0;
    for (u64 base = (u64)blockIdx.x * TILE; base < n; base += (u64)gridDim.x * TILE) {
        u64 cur[kDiffItems];
        u32 ballot[kDiffItems];
#pragma unroll
        for (int k = 0; k < kDiffItems; k++) {
            const u64 i = base + (u64)k * kDiffThreads + tid;
            bool emit = false;
            cur[k] = 0;
            if (i < n) {
                const u64 c = in[i];
                cur[k] = c;
                bool head = true, prev_del = false, differ = false;
                if (i > 0) {
                    const u64 p = in[i - 1];   // the neighbour's record: same line, served by L1
                    head = (p >> ks) != (c >> ks);
                    prev_del = rec_type(p) == kTypeDeleted;
                    differ = eff_phase(rec_phase(p)) != eff_phase(rec_phase(c));
                }
                emit = rec_type(c) == kTypeDeleted || head || prev_del || differ;
                tomb += emit && rec_type(c) == kTypeDeleted;
            }
            ballot[k] = __ballot_sync(0xffffffffu, emit);
            if (lane == 0) s_cnt[k][warp] = __popc(ballot[k]);
        }
        __syncthreads();
        if (warp == 0) {
            // exclusive scan of the 64 (item, warp) counts: two per lane
            u32 a = (&s_cnt[0][0])[2 * lane], b = (&s_cnt[0][0])[2 * lane + 1];
            u32 sum = a + b, inc = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                u32 t = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += t;
            }
            u32 total = __shfl_sync(0xffffffffu, inc, 31);
            u32 gbase = 0;
            if (lane == 0 && total) gbase = (u32)atomicAdd(&ctr->n_emitted, (u64)total);
            gbase = __shfl_sync(0xffffffffu, gbase, 0);
            (&s_cnt[0][0])[2 * lane] = inc - sum;
            (&s_cnt[0][0])[2 * lane + 1] = inc - sum + a;
            if (lane == 0) s_base = gbase;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kDiffItems; k++) {
            if ((ballot[k] >> lane) & 1u)
                out[(u64)s_base + s_cnt[k][warp] + __popc(ballot[k] & lt)] = cur[k];
        }
        __syncthreads();
    }
    tomb = __reduce_add_sync(0xffffffffu, tomb);
    if (lane == 0 && tomb) atomicAdd(&ctr->n_tombstones, (u64)tomb);
}

// ---- payload columns --------------------------------------------------------
// Emitted records, already in replay order, -> the columns a sink needs.  Four records per
// thread so that every store is a 16-byte vector (or a 4-byte word for the two byte columns).
struct AssembleArgs {
    const u64* rec; const u64* n_ptr;
    const u64* rv; const u32* ns;
    RecLayout L;
    u32* out_idx; u64* out_rv; u64* out_key; u32* out_ns; u8* out_phase; u8* out_type;
};

__global__ void __launch_bounds__(256) assemble_kernel(AssembleArgs a) {
    const u64 n = *a.n_ptr;
    const u64 stride = (u64)gridDim.x * blockDim.x * 4;
    for (u64 i = ((u64)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < n; i += stride) {
        const ulonglong2* rp = reinterpret_cast<const ulonglong2*>(a.rec + i);
        ulonglong2 r01 = rp[0], r23 = rp[1];      // padded buffers: reading past n is harmless
        u64 r[4] = {r01.x, r01.y, r23.x, r23.y};
        u32 idx[4], ns[4], ph = 0, ty = 0;
        u64 rv[4], key[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const bool ok = i + k < n;
            idx[k] = ok ? (u32)rec_idx(a.L, r[k]) : 0;
            rv[k] = ok ? __ldg(a.rv + idx[k]) : 0;
            ns[k] = ok ? __ldg(a.ns + idx[k]) : 0;
            key[k] = ok ? rec_key(a.L, r[k]) : 0;
            ph |= (ok ? rec_phase(r[k]) : 0u) << (8 * k);
            ty |= (ok ? rec_type(r[k]) : 0u) << (8 * k);
        }
        *reinterpret_cast<uint4*>(a.out_idx + i) = make_uint4(idx[0], idx[1], idx[2], idx[3]);
        *reinterpret_cast<uint4*>(a.out_ns + i) = make_uint4(ns[0], ns[1], ns[2], ns[3]);
        reinterpret_cast<ulonglong2*>(a.out_rv + i)[0] = make_ulonglong2(rv[0], rv[1]);
        reinterpret_cast<ulonglong2*>(a.out_rv + i)[1] = make_ulonglong2(rv[2], rv[3]);
        reinterpret_cast<ulonglong2*>(a.out_key + i)[0] = make_ulonglong2(key[0], key[1]);
        reinterpret_cast<ulonglong2*>(a.out_key + i)[1] = make_ulonglong2(key[2], key[3]);
        *reinterpret_cast<u32*>(a.out_phase + i) = ph;
        *reinterpret_cast<u32*>(a.out_type + i) = ty;
    }
}

// ---- digest of the emitted records (parity tool, not part of the timed path) ----
__global__ void __launch_bounds__(256)
digest_kernel(const u32* __restrict__ out_idx, const u8* __restrict__ out_phase, const u8* __restrict__ out_type,
              const u64* __restrict__ uid_hi, const u64* __restrict__ uid_lo, const u64* __restrict__ rv,
              const u32* __restrict__ ns, Counters* __restrict__ ctr) {
    const u64 n = ctr->n_emitted;
    u64 acc = 0;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        u32 j = out_idx[i];
        acc += rec_digest(i, rv[j], uid_hi[j], uid_lo[j], ns[j], out_phase[i], out_type[i]);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0 && acc) atomicAdd(&ctr->digest, acc);
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
