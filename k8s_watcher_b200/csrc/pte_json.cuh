// Raw watch-log ingestion: newline-delimited JSON watch events -> replay-log columns, on the device.
//
// A line is what the Kubernetes API streams for `?watch=true` and what the reference's loop receives as
// event['type'] / event['object'] (watcher/pod_watcher.py:264-268):
//     {"type":"MODIFIED","object":{"metadata":{"uid":"...","namespace":"...","resourceVersion":"...",...},
//                                 "spec":{...},"status":{"phase":"Running",...}}}
// Keys may come in any order, with any whitespace; everything else in the object is skipped with a
// nesting- and string-aware scan, so look-alike keys deeper down (ownerReferences[].uid, annotations that
// contain the text "phase") are never mistaken for the five fields the path reads (:162-167, :216-217).
// Namespace strings (and phase strings outside the five well-known ones) are interned EXACTLY in a device
// string dictionary: candidates are compared byte by byte against the first occurrence.
#pragma once
#include "pte_common.cuh"
#include "pte_json_scan.h"

enum : u32 { kJsonOk = 0, kJsonMalformed = 1, kJsonBadType = 2, kJsonBadUid = 3, kJsonBadRv = 4, kJsonEscapedField = 5,
             kJsonDictFull = 6, kJsonTooManyPhases = 7 };

// ---- the scanner: pte_json_scan.h (plain C++, shared with the CPU test harness) ----------------------------
typedef ptejson::Span JsonSpan;
typedef ptejson::Fields JsonFields;

// the text in global memory: bytes by offset, 32-byte blocks by aligned offset (two 16-byte loads), eight bytes from any offset
struct GlobalText {
    const char* base;
    __device__ __forceinline__ char operator[](u32 i) const { return base[i]; }
    __device__ __forceinline__ void load32(u32 blk, u32 (&w)[8]) const {
        const uint4* g = reinterpret_cast<const uint4*>(base + blk);
        const uint4 a = __ldg(g), b = __ldg(g + 1);
        w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
    }
    __device__ __forceinline__ u64 load8(u32 s) const {
        const u64* g = reinterpret_cast<const u64*>(base + (s & ~7u));
        const u64 a = __ldg(g), b = __ldg(g + 1);
        const u32 sh = (s & 7u) * 8u;
        return sh ? (a >> sh) | (b << (64u - sh)) : a;
    }
};
// a few decoded bytes in local memory, for the rare field that is spelled with escapes
struct LocalText {
    const char* base;
    __device__ __forceinline__ char operator[](u32 i) const { return base[i]; }
};
template <typename T>
__device__ __forceinline__ bool has_backslash(const T& t, u32 s, u32 len) {
    bool any = false;
    for (u32 k = 0; k < len; k++) any = any || t[s + k] == '\\';
    return any;
}

// ---- exact string dictionary -----------------------------------------------------------------------
// Open addressing; a slot is claimed with a 64-bit (hash, length) key and remembers the smallest text
// offset at which its string occurs.  A candidate with an equal key is compared byte by byte with that
// occurrence, so two different strings never share a slot.
struct StrDict { u64* keys; u32* first_off; u32 mask; };
static constexpr u32 kNoOffset = 0xFFFFFFFFu;

// [s, s + raw_len) is the candidate's string body in the text.  Strings are the same when the bytes they STAND FOR are
// the same: "ab" and "a\u0062" share a slot.  The hash and the stored length are those of the decoded bytes; the common
// case (no backslash on either side) never decodes anything.
template <typename T>
__device__ __forceinline__ u32 dict_intern(const StrDict& d, const T& t, u32 s, u32 raw_len, u32& bad) {
    u64 h = 0xCBF29CE484222325ull;
    bool esc = false;
    for (u32 i = 0; i < raw_len; i++) { const char c = t[s + i]; esc = esc || c == '\\'; h = (h ^ (u8)c) * 0x100000001B3ull; }
    u32 len = raw_len;
    if (esc) {
        ptejson::Unescape<T> u(t, s);
        h = 0xCBF29CE484222325ull; len = 0;
        while ((u.pos < s + raw_len || u.n_queue) && !u.bad) { h = (h ^ u.next()) * 0x100000001B3ull; len++; }
        if (u.bad || u.pos > s + raw_len) { bad = kJsonEscapedField; return 0; }
    }
    h = mix64(h);
    const u64 key = (h & ~0xFFFFFull) | (len & 0xFFFFFu) | (1ull << 63);
    u32 slot = (u32)(h >> 24) & d.mask;
    for (u32 probe = 0; probe <= d.mask; probe++) {
        u64 k = *reinterpret_cast<volatile u64*>(d.keys + slot);
        if (k == 0) {
            const u64 old = atomicCAS(d.keys + slot, 0ull, key);
            if (old == 0) { atomicMin(d.first_off + slot, s); return slot; }
            k = old;
        }
        if (k == key) {
            u32 f;
            while ((f = *reinterpret_cast<volatile u32*>(d.first_off + slot)) == kNoOffset) __nanosleep(20);   // claimer is publishing
            bool same = true, slow = esc;
            for (u32 i = 0; !slow && i < len && same; i++) { const char c = t[f + i]; slow = c == '\\'; same = c == t[s + i]; }
            if (slow) {
                ptejson::Unescape<T> x(t, f), y(t, s);
                same = true;
                for (u32 i = 0; i < len && same; i++) same = x.next() == y.next();
            }
            if (same) { if (s < f) atomicMin(d.first_off + slot, s); return slot; }     // hot slots: almost never an atomic
        }
        slot = (slot + 1) & d.mask;
    }
    bad = kJsonDictFull;
    return 0;
}

// ---- kernels ------------------------------------------------------------------------------------------
static constexpr int kJsonChunk = 4096;     // bytes of text per block in the newline kernels

// newline mask of the 16 bytes at t + base (bit k = byte k); bytes at or behind len are not looked at.
// The text buffer is 256-byte aligned and zero-padded behind len (pte_ingest_json).
__device__ __forceinline__ u32 newline_mask16(const char* __restrict__ t, u64 len, u64 base) {
    if (base >= len) return 0;
    const uint4 v = __ldcs(reinterpret_cast<const uint4*>(t + base));
    return ptejson::eq4(v.x, 0x0A0A0A0Au) | (ptejson::eq4(v.y, 0x0A0A0A0Au) << 4) | (ptejson::eq4(v.z, 0x0A0A0A0Au) << 8) |
           (ptejson::eq4(v.w, 0x0A0A0A0Au) << 12);
}

__global__ void __launch_bounds__(256) json_count_newlines_kernel(const char* __restrict__ t, u64 len, u32* __restrict__ chunk_count) {
    __shared__ u32 s_w[8];
    const u64 n_chunks = (len + kJsonChunk - 1) / kJsonChunk;
    for (u64 c = blockIdx.x; c < n_chunks; c += gridDim.x) {
        u32 cnt = __popc(newline_mask16(t, len, c * kJsonChunk + threadIdx.x * 16));
        cnt = __reduce_add_sync(0xffffffffu, cnt);
        if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = cnt;
        __syncthreads();
        if (threadIdx.x == 0) { u32 v = 0; for (int k = 0; k < 8; k++) v += s_w[k]; chunk_count[c] = v; }
        __syncthreads();
    }
}

// newline j of the text is at nl[j]; chunk_off = exclusive scan of chunk_count
__global__ void __launch_bounds__(256)
json_list_newlines_kernel(const char* __restrict__ t, u64 len, const u64* __restrict__ chunk_off, u64* __restrict__ nl) {
    __shared__ u32 s_cnt[8];
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const u64 n_chunks = (len + kJsonChunk - 1) / kJsonChunk;
    for (u64 c = blockIdx.x; c < n_chunks; c += gridDim.x) {
        // each warp owns 512 consecutive bytes of the chunk, each lane 16 consecutive bytes of those
        const u64 base = c * kJsonChunk + threadIdx.x * 16;
        u32 bits = newline_mask16(t, len, base);
        const u32 mine = __popc(bits);
        u32 inc = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { u32 v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
        if (lane == 31) s_cnt[warp] = inc;
        __syncthreads();
        u64 dst = chunk_off[c] + inc - mine;
        for (u32 w = 0; w < warp; w++) dst += s_cnt[w];
        while (bits) { const u32 k = __ffs(bits) - 1; bits &= bits - 1; nl[dst++] = base + k; }
        __syncthreads();
    }
}

struct JsonArgs {
    const char* text; u64 len;
    const u64* nl; u64 n_lines;            // line i = [i ? nl[i-1] + 1 : 0, i < n_nl ? nl[i] : len)
    u64 n_nl;
    u64* uid_hi; u64* uid_lo; u64* rv; u32* ns; u8* phase; u8* type;
    StrDict ns_dict, phase_dict, uid_dict;
    u32* err;          // [0] worst error code, [1] (via atomicMin) first bad line + 1
    int arrival_order; u64 arrival_base;   // stream mode: row i gets the order key arrival_base + i + 1, resourceVersion is not read
};

// ---- the five fields -> codes -------------------------------------------------------------------------
// (templates over where the bytes come from: the text, or a few decoded bytes in local memory)
template <typename A>
__device__ __forceinline__ bool parse_type(const A& t, u32 s, u32 len, u32& type) {
    // event['type'] (pod_watcher.py:267): ADDED / MODIFIED / DELETED, anything else is rejected
    const char c0 = len ? t[s] : '\0';
    const char* name;
    if (len == 5 && c0 == 'A') { type = kTypeAdded; name = "ADDED"; }
    else if (len == 8 && c0 == 'M') { type = kTypeModified; name = "MODIFIED"; }
    else if (len == 7 && c0 == 'D') { type = kTypeDeleted; name = "DELETED"; }
    else return false;
    bool ok = true;
    for (u32 k = 1; k < len; k++) ok = ok && t[s + k] == name[k];
    return ok;
}
// metadata.uid (:164) in the form Kubernetes emits -- lower-case hex, 8-4-4-4-12 -- is its own 128-bit key.  Whatever else
// arrives (the reference takes any string) is interned exactly in a string dictionary and gets a key of the reserved block
// hi == kUidReservedHi; so do the canonical strings whose value lies in that block or is zero.  Two different uid strings
// never share a key.  (codes.UidInterner is the host's form of the same rule.)
static constexpr u64 kUidReservedHi = 0xFFFFFFFFFFFFFFFFull;
template <typename A>
__device__ __forceinline__ bool parse_canonical_uid(const A& t, u32 s, u32 len, u64& hi, u64& lo) {
    if (len != 36) return false;
    hi = lo = 0;
    for (u32 k = 0; k < 36; k++) {
        const char c = t[s + k];
        if (k == 8 || k == 13 || k == 18 || k == 23) { if (c != '-') return false; continue; }
        int v;
        if (c >= '0' && c <= '9') v = c - '0'; else if (c >= 'a' && c <= 'f') v = c - 'a' + 10; else return false;
        const u32 nd = k - (k > 8) - (k > 13) - (k > 18) - (k > 23);
        if (nd < 16) hi = (hi << 4) | (u64)v; else lo = (lo << 4) | (u64)v;
    }
    return hi != kUidReservedHi && (hi | lo) != 0;
}
template <typename A>
__device__ __forceinline__ bool parse_rv(const A& t, u32 s, u32 len, u64& rv) {
    if (len == 0 || len > 20) return false;
    rv = 0;
    for (u32 k = 0; k < len; k++) {
        const char c = t[s + k];
        if (c < '0' || c > '9') return false;
        rv = rv * 10 + (u64)(c - '0');
    }
    return true;
}
template <typename A>
__device__ __forceinline__ u32 known_phase(const A& t, u32 s, u32 len) {
    const char* known[5] = {"Pending", "Running", "Succeeded", "Failed", "Unknown"};
    const u32 klen[5] = {7, 7, 9, 6, 7};
    for (u32 c = 0; c < 5; c++) {
        if (len != klen[c]) continue;
        bool same = true;
        for (u32 k = 0; k < klen[c] && same; k++) same = t[s + k] == known[c][k];
        if (same) return c;
    }
    return 0xFFu;
}
// decoded bytes of a short string spelled with escapes.  0: n bytes in buf, 1: longer than cap, 2: broken escape
static constexpr u32 kJsonSmall = 40;
template <typename T>
__device__ __noinline__ u32 decode_small(const T& t, u32 s, u32 raw_len, char* buf, u32& n) {
    ptejson::Unescape<T> u(t, s);
    n = 0;
    while (u.pos < s + raw_len || u.n_queue) {
        const u32 b = u.next();
        if (u.bad) return 2;
        if (n == kJsonSmall) return 1;
        buf[n++] = (char)b;
    }
    return u.pos > s + raw_len ? 2u : 0u;
}

// One line [p, end) of the text -> row i of the columns.
static constexpr int kJsonThreads = 128;
// the scanner's candidate queue of one thread: word k of entry e at base[(4 e + k) * blockDim + thread] (no bank conflicts)
struct SmemQueue {
    u32* base;
    __device__ __forceinline__ void put(u32 e, u32 k, u32 v) { base[(4 * e + k) * kJsonThreads] = v; }
    __device__ __forceinline__ u32 get(u32 e, u32 k) const { return base[(4 * e + k) * kJsonThreads]; }
};

template <typename T>
__device__ __forceinline__ void json_line(const JsonArgs& a, const T& t, u64 i, u32 p, u32 end, SmemQueue& q) {
    using namespace ptejson;
    u32 bad = kJsonOk;
    JsonFields f;
    u64 hi = 0, lo = 0, rv = i + 1;
    u32 type = 0, phase = kPhNoStatus, ns = 0;
    char buf[kJsonSmall];
    u32 n = 0;
    if (!scan_line(t, p, end, f, q)) bad = kJsonMalformed;
    if (!bad) {
        const u32 s = span_start(f.type), len = span_len(f.type);
        bool ok = span_present(f.type);
        if (ok && has_backslash(t, s, len)) {
            const u32 r = decode_small(t, s, len, buf, n);
            if (r == 2) bad = kJsonEscapedField;
            ok = r == 0 && parse_type(LocalText{buf}, 0, n, type);
        } else ok = ok && parse_type(t, s, len, type);
        if (!bad && !ok) bad = kJsonBadType;
    }
    if (!bad) {
        const u32 s = span_start(f.uid), len = span_len(f.uid);
        if (!span_present(f.uid)) bad = kJsonBadUid;               // no uid string at all
        else {
            bool direct;
            if (has_backslash(t, s, len)) {
                const u32 r = decode_small(t, s, len, buf, n);
                if (r == 2) bad = kJsonEscapedField;
                direct = r == 0 && parse_canonical_uid(LocalText{buf}, 0, n, hi, lo);
            } else direct = parse_canonical_uid(t, s, len, hi, lo);
            if (!bad && !direct) {
                hi = kUidReservedHi;
                lo = dict_intern(a.uid_dict, t, s, len, bad);      // slot for now; json_remap_kernel turns it into the key
                atomicAdd(a.err + 3, 1u);
            }
        }
    }
    if (a.arrival_order) rv = a.arrival_base + i + 1;      // the reference's loop knows no other order (it never reads resourceVersion)
    else if (!bad && span_present(f.rv)) {
        const u32 s = span_start(f.rv), len = span_len(f.rv);
        bool ok;
        if (has_backslash(t, s, len)) {
            const u32 r = decode_small(t, s, len, buf, n);
            if (r == 2) bad = kJsonEscapedField;
            ok = r == 0 && parse_rv(LocalText{buf}, 0, n, rv);
        } else ok = parse_rv(t, s, len, rv);
        if (!bad && !ok) bad = kJsonBadRv;
    }
    if (!bad) {
        // metadata.namespace (:163): any string; a pod without one is in the namespace "" like every other such pod
        ns = dict_intern(a.ns_dict, t, span_present(f.ns) ? span_start(f.ns) : p, span_present(f.ns) ? span_len(f.ns) : 0u, bad);
        // status.phase (:167): NoStatus when pod.status is falsy, Null when there is no phase string
        if (!f.has_status) phase = kPhNoStatus;
        else if (!span_present(f.phase)) phase = kPhNull;
        else {
            const u32 s = span_start(f.phase), len = span_len(f.phase);
            if (has_backslash(t, s, len)) phase = decode_small(t, s, len, buf, n) == 0 ? known_phase(LocalText{buf}, 0, n) : 0xFFu;
            else phase = known_phase(t, s, len);
            if (phase == 0xFFu) phase = 0x80u | dict_intern(a.phase_dict, t, s, len, bad);   // remapped later
        }
    }
    if (bad) { atomicMax(a.err, bad); atomicMin(a.err + 1, (u32)min(i, (u64)0xFFFFFFFEu)); }
    a.uid_hi[i] = hi; a.uid_lo[i] = lo; a.rv[i] = rv; a.ns[i] = ns; a.phase[i] = (u8)phase; a.type[i] = (u8)type;
}

// One thread per line, 32 consecutive lines per warp (handed out by ticket, so long and short lines even out); the
// text comes straight from global memory in 32-byte blocks, the next block in flight while the current one is scanned.
#ifndef PTE_JSON_BLOCKS
#define PTE_JSON_BLOCKS 6      // 8 (64 registers, spills) measured slower: 2.36 vs 2.30 ms per GB
#endif

__global__ void __launch_bounds__(kJsonThreads, PTE_JSON_BLOCKS) json_parse_kernel(JsonArgs a, u32* __restrict__ ticket) {
    __shared__ u32 s_queue[ptejson::kQueueCap * 4 * kJsonThreads];
    const u32 lane = threadIdx.x & 31;
    const GlobalText t{a.text};
    SmemQueue q{s_queue + threadIdx.x};
    for (;;) {
        u64 g0 = 0;
        if (lane == 0) g0 = (u64)atomicAdd(ticket, 1u) * 32;
        g0 = __shfl_sync(0xffffffffu, g0, 0);
        if (g0 >= a.n_lines) break;
        const u64 li = g0 + lane;
        if (li < a.n_lines) {
            const u32 p = li ? (u32)a.nl[li - 1] + 1u : 0u;
            const u32 end = li < a.n_nl ? (u32)a.nl[li] : (u32)a.len;
            json_line(a, t, li, p, end, q);
        }
        __syncwarp();
    }
}

// dictionary slots -> dense ids in order of first occurrence (remap tables built by the host)
__global__ void __launch_bounds__(256)
json_remap_kernel(u32* __restrict__ ns, u8* __restrict__ phase, u64 n, const u32* __restrict__ ns_remap, const u8* __restrict__ phase_remap,
                  const u64* __restrict__ uid_hi, u64* __restrict__ uid_lo, const u32* __restrict__ uid_remap) {
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        ns[i] = ns_remap[ns[i]];
        if (uid_remap && uid_hi[i] == kUidReservedHi) uid_lo[i] = (u64)uid_remap[uid_lo[i]] + 1;      // reserved keys start at 1
        const u8 ph = phase[i];
        if (ph & 0x80u) phase[i] = phase_remap[ph & 0x7Fu];
    }
}
