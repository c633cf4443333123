// Shared types, the 64-bit sort-record layout and the synthetic-log arithmetic
// (host and device use the same inline functions, so the device generator and
// pte_generate_host cannot drift apart).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

typedef unsigned long long u64;
typedef unsigned int u32;
typedef unsigned char u8;

#define PTE_HD __host__ __device__ __forceinline__

PTE_HD u64 mix64(u64 z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

PTE_HD u64 unxorshift64(u64 x, int s) {
    u64 r = x;
    for (int k = s; k < 64; k += s) r = x ^ (r >> s);
    return r;
}
PTE_HD u64 unmix64(u64 z) {
    z = unxorshift64(z, 31) * 0x319642B2D24D8EC3ull;
    z = unxorshift64(z, 27) * 0x96DE1B173F119089ull;
    return unxorshift64(z, 30);
}

PTE_HD u64 mulhi64(u64 a, u64 b) {
#ifdef __CUDA_ARCH__
    return __umul64hi(a, b);
#else
    return (u64)(((unsigned __int128)a * b) >> 64);
#endif
}

static constexpr u64 kUidA = 0x6B38735F77617463ull;
static constexpr u64 kUidB = 0x5F62323030706F64ull;

// ---- event codes (mirror include/pte.h) ----
enum : u32 { kTypeAdded = 0, kTypeModified = 1, kTypeDeleted = 2 };
enum : u32 { kPhSucceeded = 2, kPhFailed = 3, kPhUnknown = 4, kPhNull = 5, kPhNoStatus = 6, kPhMax = 62 };

// Phase the payload carries: a pod without status reports "Unknown" (pod_watcher.py:167).
PTE_HD u32 eff_phase(u32 ph) { return ph == kPhNoStatus ? (u32)kPhUnknown : ph; }

// ---- 64-bit sort record -----------------------------------------------------
//   [0,6) phase   [6,8) type   [8, 8+IB) arrival index   [8+IB, 8+IB+SB) interned key (table slot)
// A stable LSD sort on the key field of records laid out in arrival order is a
// sort on (key, rv) for an rv-ordered log, which is what the diff stage needs.
struct RecLayout {
    u32 idx_bits;   // IB
    u32 key_bits;   // SB
    PTE_HD u32 key_shift() const { return 8 + idx_bits; }
    PTE_HD u64 idx_mask() const { return (1ull << idx_bits) - 1; }
};
PTE_HD u64 rec_pack(const RecLayout& L, u64 key, u64 idx, u32 type, u32 phase) {
    return (key << L.key_shift()) | (idx << 8) | ((u64)type << 6) | phase;
}
PTE_HD u32 rec_phase(u64 r) { return (u32)(r & 63); }
PTE_HD u32 rec_type(u64 r) { return (u32)((r >> 6) & 3); }
PTE_HD u64 rec_idx(const RecLayout& L, u64 r) { return (r >> 8) & L.idx_mask(); }
PTE_HD u64 rec_key(const RecLayout& L, u64 r) { return r >> L.key_shift(); }

// ---- order-sensitive digest (same arithmetic as oracle/oracle.c rec_digest) ----
PTE_HD u64 rec_digest(u64 pos, u64 rv, u64 hi, u64 lo, u32 ns, u32 phase, u32 type) {
    u64 small = ((u64)ns << 16) | ((u64)phase << 8) | type;
    u64 h = mix64(lo ^ small);
    h = mix64(hi ^ h);
    h = mix64(rv ^ h);
    return mix64(((pos + 1) * 0x9E3779B97F4A7C15ull) ^ h);
}

// ---- synthetic replay log: counter-based part (spec in include/pte.h) ----
struct GenParams {
    u64 seed, n_events, n_pods;
    u32 dist, p_del_q24, p_nostatus_q24, p_null_q24, shard, n_shards;
};
struct GenEvent {
    u64 q, pod, rv, uid_hi, uid_lo;
    u32 ns, phase, draw;
};
PTE_HD u32 bitlen64(u64 x) {
    u32 n = 0;
    while (x) { n++; x >>= 1; }
    return n;
}
PTE_HD GenEvent gen_event(const GenParams& g, u64 j) {
    GenEvent o;
    u64 e = j * g.n_shards + g.shard;
    u64 r1 = mix64(g.seed ^ mix64(e + 0x9E3779B97F4A7C15ull));
    u64 r2 = mix64(r1 + 0xD1B54A32D192ED03ull);
    u64 r3 = mix64(r2 + 0x8CB92BA72F3D8DD7ull);
    u64 r4 = mix64(r3 + 0x2545F4914F6CDD1Dull);
    u64 P = g.n_pods;
    if (g.dist == 0) {
        o.q = mulhi64(r1, P);
    } else {
        u64 oct = mulhi64(r1, (u64)bitlen64(P));
        u64 width = 1ull << oct;
        u64 rank = (width - 1) + (r2 & (width - 1));
        if (rank >= P) rank %= P;
        o.q = (rank * 2654435761ull + 40503ull) % P;
    }
    o.pod = o.q * g.n_shards + g.shard;
    o.rv = e + 1;
    u32 ph = (u32)(((r3 & 0xFFFFFFFFull) * 5) >> 32);
    if ((u32)(r4 & 0xFFFFFF) < g.p_nostatus_q24) ph = kPhNoStatus;
    else if ((u32)((r4 >> 24) & 0xFFFFFF) < g.p_null_q24) ph = kPhNull;
    o.phase = ph;
    o.draw = (u32)(r3 >> 40) < g.p_del_q24;
    o.uid_hi = mix64(o.pod ^ kUidA);
    o.uid_lo = mix64(~o.pod ^ kUidB);
    o.ns = (u32)(o.pod % 5);
    return o;
}
// type from (has predecessor, predecessor's draw, own draw)
PTE_HD u32 gen_type(bool has_prev, u32 prev_draw, u32 draw) {
    if (!has_prev || prev_draw) return kTypeAdded;
    return draw ? kTypeDeleted : kTypeModified;
}

PTE_HD u32 shard_of_uid(u64 uid_hi, u32 n_shards) { return (u32)((unmix64(uid_hi) ^ kUidA) % n_shards); }
