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

// the text in global memory: bytes by offset, 32-byte blocks by aligned offset (two 16-byte loads)
struct GlobalText {
    const char* base;
    __device__ __forceinline__ char operator[](u64 i) const { return base[i]; }
    __device__ __forceinline__ void load32(u64 blk, u32 (&w)[8]) const {
        const uint4* g = reinterpret_cast<const uint4*>(base + blk);
        const uint4 a = __ldg(g), b = __ldg(g + 1);
        w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
    }
};
template <typename T>
__device__ __forceinline__ bool span_has_backslash(const T& t, const JsonSpan& sp) {
    bool any = false;
    for (u32 k = 0; sp.present && k < sp.len; k++) any = any || t[sp.s + k] == '\\';
    return any;
}

// ---- exact string dictionary -----------------------------------------------------------------------
// Open addressing; a slot is claimed with a 64-bit (hash, length) key and remembers the smallest text
// offset at which its string occurs.  A candidate with an equal key is compared byte by byte with that
// occurrence, so two different strings never share a slot.
struct StrDict { u64* keys; u32* first_off; u32 mask; };
static constexpr u32 kNoOffset = 0xFFFFFFFFu;

// t: where the candidate's bytes are read from (global text or a shared-memory window of it); g: the global text,
// against which earlier occurrences are compared.  s is the candidate's offset in the global text.
template <typename T>
__device__ __forceinline__ u32 dict_intern(const StrDict& d, const char* g, const T& t, u64 s, u32 len, u32* err) {
    u64 h = 0xCBF29CE484222325ull;
    for (u32 i = 0; i < len; i++) h = (h ^ (u8)t[s + i]) * 0x100000001B3ull;
    h = mix64(h);
    const u64 key = (h & ~0xFFFFFull) | (len & 0xFFFFFu) | (1ull << 63);
    u32 slot = (u32)(h >> 24) & d.mask;
    for (u32 probe = 0; probe <= d.mask; probe++) {
        u64 k = *reinterpret_cast<volatile u64*>(d.keys + slot);
        if (k == 0) {
            const u64 old = atomicCAS(d.keys + slot, 0ull, key);
            if (old == 0) { atomicMin(d.first_off + slot, (u32)s); return slot; }
            k = old;
        }
        if (k == key) {
            u32 f;
            while ((f = *reinterpret_cast<volatile u32*>(d.first_off + slot)) == kNoOffset) __nanosleep(20);   // claimer is publishing
            bool same = true;
            for (u32 i = 0; i < len && same; i++) same = g[f + i] == t[s + i];
            if (same) { if ((u32)s < f) atomicMin(d.first_off + slot, (u32)s); return slot; }     // hot slots: almost never an atomic
        }
        slot = (slot + 1) & d.mask;
    }
    atomicMax(err, (u32)kJsonDictFull);
    return 0;
}

// ---- kernels ------------------------------------------------------------------------------------------
static constexpr int kJsonChunk = 4096;     // bytes of text per block in the newline kernels

__global__ void __launch_bounds__(256) json_count_newlines_kernel(const char* __restrict__ t, u64 len, u32* __restrict__ chunk_count) {
    __shared__ u32 s_w[8];
    const u64 n_chunks = (len + kJsonChunk - 1) / kJsonChunk;
    for (u64 c = blockIdx.x; c < n_chunks; c += gridDim.x) {
        u32 cnt = 0;
        for (u32 k = threadIdx.x; k < kJsonChunk; k += 256) {
            const u64 i = c * kJsonChunk + k;
            cnt += i < len && t[i] == '\n';
        }
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
        const u64 base = c * kJsonChunk + warp * 512 + lane * 16;
        u32 bits = 0;
        for (u32 k = 0; k < 16; k++) bits |= (u32)(base + k < len && t[base + k] == '\n') << k;
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
    StrDict ns_dict, phase_dict;
    u32* err;          // [0] worst error code, [1] (via atomicMin) first bad line + 1
};

__device__ __forceinline__ int hex_val(char c) {
    if (c >= '0' && c <= '9') return c - '0';
    if (c >= 'a' && c <= 'f') return c - 'a' + 10;
    if (c >= 'A' && c <= 'F') return c - 'A' + 10;
    return -1;
}

// One line [p, end) of the text -> row i of the columns.  t gives access to the line's bytes by their offset in the text.
template <typename T>
__device__ __forceinline__ void json_line(const JsonArgs& a, const T& t, u64 i, u64 p, u64 end) {
    {
        u32 bad = kJsonOk;
        JsonFields f;
        u64 hi = 0, lo = 0, rv = i + 1;
        u32 type = 0, phase = kPhNoStatus, ns = 0;
        if (!ptejson::scan_line(t, p, end, f)) bad = kJsonMalformed;
        if (!bad && (span_has_backslash(t, f.type) || span_has_backslash(t, f.uid) || span_has_backslash(t, f.ns) ||
                     span_has_backslash(t, f.rv) || span_has_backslash(t, f.phase))) bad = kJsonEscapedField;
        if (!bad) {
            // event['type'] (pod_watcher.py:267): ADDED / MODIFIED / DELETED, anything else is rejected
            const char c0 = f.type.present && f.type.len ? t[f.type.s] : 0;
            if (f.type.present && f.type.len == 5 && c0 == 'A') type = kTypeAdded;
            else if (f.type.present && f.type.len == 8 && c0 == 'M') type = kTypeModified;
            else if (f.type.present && f.type.len == 7 && c0 == 'D') type = kTypeDeleted;
            else bad = kJsonBadType;
            const char* names[3] = {"ADDED", "MODIFIED", "DELETED"};
            if (!bad) for (u32 k = 0; k < f.type.len; k++) if (t[f.type.s + k] != names[type][k]) bad = kJsonBadType;
        }
        if (!bad) {
            // metadata.uid (:164): 32 hex digits, dashes ignored
            u32 nd = 0;
            for (u32 k = 0; f.uid.present && k < f.uid.len; k++) {
                const char c = t[f.uid.s + k];
                if (c == '-') continue;
                const int v = hex_val(c);
                if (v < 0 || nd >= 32) { nd = 99; break; }
                if (nd < 16) hi = (hi << 4) | (u64)v; else lo = (lo << 4) | (u64)v;
                nd++;
            }
            if (nd != 32 || (hi | lo) == 0) bad = kJsonBadUid;
        }
        if (!bad && f.rv.present) {
            rv = 0;
            if (f.rv.len == 0 || f.rv.len > 20) bad = kJsonBadRv;
            for (u32 k = 0; !bad && k < f.rv.len; k++) {
                const char c = t[f.rv.s + k];
                if (c < '0' || c > '9') bad = kJsonBadRv; else rv = rv * 10 + (u64)(c - '0');
            }
        }
        if (!bad) {
            ns = dict_intern(a.ns_dict, a.text, t, f.ns.present ? f.ns.s : p, f.ns.present ? f.ns.len : 0, a.err);
            if (!f.has_status) phase = kPhNoStatus;
            else if (!f.phase.present) phase = kPhNull;
            else {
                const char* known[5] = {"Pending", "Running", "Succeeded", "Failed", "Unknown"};
                const u32 klen[5] = {7, 7, 9, 6, 7};
                phase = 0xFFu;
                for (u32 c = 0; c < 5 && phase == 0xFFu; c++) {
                    if (f.phase.len != klen[c]) continue;
                    bool same = true;
                    for (u32 k = 0; k < klen[c] && same; k++) same = t[f.phase.s + k] == known[c][k];
                    if (same) phase = c;
                }
                if (phase == 0xFFu) phase = 0x80u | dict_intern(a.phase_dict, a.text, t, f.phase.s, f.phase.len, a.err);   // remapped later
            }
        }
        if (bad) { atomicMax(a.err, bad); atomicMin(a.err + 1, (u32)min(i, (u64)0xFFFFFFFEu));
        }
        a.uid_hi[i] = hi; a.uid_lo[i] = lo; a.rv[i] = rv; a.ns[i] = ns; a.phase[i] = (u8)phase; a.type[i] = (u8)type;
    }
}

// One thread per line, 32 consecutive lines per warp (handed out by ticket, so long and short lines even out); the
// text comes straight from global memory in 32-byte blocks, the next block in flight while the current one is scanned.
static constexpr int kJsonThreads = 128;

__global__ void __launch_bounds__(kJsonThreads) json_parse_kernel(JsonArgs a, u32* __restrict__ ticket) {
    const u32 lane = threadIdx.x & 31;
    const GlobalText t{a.text};
    for (;;) {
        u64 g0 = 0;
        if (lane == 0) g0 = (u64)atomicAdd(ticket, 1u) * 32;
        g0 = __shfl_sync(0xffffffffu, g0, 0);
        if (g0 >= a.n_lines) break;
        const u64 li = g0 + lane;
        if (li < a.n_lines) {
            const u64 p = li ? a.nl[li - 1] + 1 : 0;
            const u64 end = li < a.n_nl ? a.nl[li] : a.len;
            json_line(a, t, li, p, end);
        }
        __syncwarp();
    }
}

// dictionary slots -> dense ids in order of first occurrence (remap tables built by the host)
__global__ void __launch_bounds__(256)
json_remap_kernel(u32* __restrict__ ns, u8* __restrict__ phase, u64 n, const u32* __restrict__ ns_remap, const u8* __restrict__ phase_remap) {
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        ns[i] = ns_remap[ns[i]];
        const u8 ph = phase[i];
        if (ph & 0x80u) phase[i] = phase_remap[ph & 0x7Fu];
    }
}
