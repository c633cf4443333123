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

enum : u32 { kJsonOk = 0, kJsonMalformed = 1, kJsonBadType = 2, kJsonBadUid = 3, kJsonBadRv = 4, kJsonEscapedField = 5,
             kJsonDictFull = 6, kJsonTooManyPhases = 7 };

struct JsonSpan { u64 s; u32 len; bool present, escaped; };

// ---- the scanner ---------------------------------------------------------------------------------------
// One pass over the bytes of a line with a handful of state variables, the same few instructions for every byte:
// threads of a warp scan 32 different lines, and a recursive-descent scanner (one code path per syntactic
// situation) had them take turns -- 300 k warp instructions per 32 lines, measured.  State: nesting depth, a bit
// per depth saying "object or array", the meaning of the containers at depths 1..3 (root / object / metadata /
// status / other), whether a string is a key, and which field the value after the last key is for.
struct JsonFields { JsonSpan type, uid, ns, rv, phase; bool has_status; };

__host__ __device__ constexpr u32 json_fnv(const char* s, u32 n) {
    u32 h = 2166136261u;
    for (u32 i = 0; i < n; i++) h = (h ^ (u32)(unsigned char)s[i]) * 16777619u;
    return h;
}
template <typename T>
__device__ __forceinline__ bool json_bytes_are(const T& t, u64 s, const char* name, u32 n) {
    for (u32 i = 0; i < n; i++) if (t[s + i] != name[i]) return false;
    return true;
}

template <typename T>
__device__ __forceinline__ bool json_scan_line(const T& t, u64 p, u64 end, JsonFields& f) {
    enum : u32 { C_OTHER = 0, C_ROOT = 1, C_OBJECT = 2, C_META = 3, C_STATUS = 4 };
    enum : u32 { W_NONE = 0, W_TYPE = 1, W_UID = 2, W_NS = 3, W_RV = 4, W_PHASE = 5, W_OBJECT = 6, W_META = 7, W_STATUS = 8 };
    const JsonSpan none = {0, 0, false, false};
    f.type = f.uid = f.ns = f.rv = f.phase = none;
    f.has_status = false;
    u32 depth = 0, ctxs = 0, want = W_NONE, h = 0;
    u64 is_obj = 0, str_start = 0;
    bool in_str = false, esc = false, str_esc = false, is_key = false, expect_key = false, bad = false, opened = false;
    u32 key_h = 0, key_len = 0;
    u64 key_s = 0;
    bool key_ok = false;
    for (; p < end; p++) {
        const char c = t[p];
        if (in_str) {
            if (esc) esc = false;
            else if (c == '\\') { esc = true; str_esc = true; }
            else if (c == '"') {
                in_str = false;
                if (is_key) { key_h = h; key_len = (u32)(p - str_start); key_s = str_start; key_ok = !str_esc; }
                else {
                    if (want >= W_TYPE && want <= W_PHASE) {
                        const JsonSpan sp = {str_start, (u32)(p - str_start), true, str_esc};
                        if (want == W_TYPE) f.type = sp; else if (want == W_UID) f.uid = sp; else if (want == W_NS) f.ns = sp;
                        else if (want == W_RV) f.rv = sp; else f.phase = sp;
                    }
                    want = W_NONE;
                }
            } else h = (h ^ (u32)(unsigned char)c) * 16777619u;
            continue;
        }
        if (c == ' ' || c == '\t' || c == '\r' || c == '\n') continue;
        if (c == '"') {
            in_str = true; str_esc = false; h = 2166136261u; str_start = p + 1;
            is_key = expect_key; expect_key = false;
        } else if (c == ':') {
            // the key just read decides what the value is for
            const u32 ctx = depth <= 7 ? (ctxs >> (4 * depth)) & 15u : C_OTHER;
            want = W_NONE;
            if (key_ok && ctx != C_OTHER) {
                u32 w = W_NONE; const char* name = ""; u32 n = 0;
                if (ctx == C_ROOT) {
                    if (key_len == 4 && key_h == json_fnv("type", 4)) { w = W_TYPE; name = "type"; n = 4; }
                    else if (key_len == 6 && key_h == json_fnv("object", 6)) { w = W_OBJECT; name = "object"; n = 6; }
                } else if (ctx == C_OBJECT) {
                    if (key_len == 8 && key_h == json_fnv("metadata", 8)) { w = W_META; name = "metadata"; n = 8; }
                    else if (key_len == 6 && key_h == json_fnv("status", 6)) { w = W_STATUS; name = "status"; n = 6; }
                } else if (ctx == C_META) {
                    if (key_len == 3 && key_h == json_fnv("uid", 3)) { w = W_UID; name = "uid"; n = 3; }
                    else if (key_len == 9 && key_h == json_fnv("namespace", 9)) { w = W_NS; name = "namespace"; n = 9; }
                    else if (key_len == 15 && key_h == json_fnv("resourceVersion", 15)) { w = W_RV; name = "resourceVersion"; n = 15; }
                } else if (ctx == C_STATUS) {
                    if (key_len == 5 && key_h == json_fnv("phase", 5)) { w = W_PHASE; name = "phase"; n = 5; }
                }
                if (w != W_NONE && json_bytes_are(t, key_s, name, n)) {      // the hash only nominates, the bytes decide
                    want = w;
                    // a value that is not a string (null, a number) leaves the field absent; a later duplicate key wins
                    if (w == W_TYPE) f.type = none; else if (w == W_UID) f.uid = none; else if (w == W_NS) f.ns = none;
                    else if (w == W_RV) f.rv = none; else if (w == W_PHASE) f.phase = none;
                }
            }
        } else if (c == '{' || c == '[') {
            u32 nc = C_OTHER;
            if (c == '{') {
                if (depth == 0) nc = C_ROOT;
                else if (want == W_OBJECT) nc = C_OBJECT;
                else if (want == W_META) nc = C_META;
                else if (want == W_STATUS) { nc = C_STATUS; f.has_status = true; }      // pod.status is truthy (pod_watcher.py:167,209)
            }
            opened = true;
            depth++;
            if (depth <= 7) ctxs = (ctxs & ~(15u << (4 * depth))) | (nc << (4 * depth));
            if (depth < 64) is_obj = c == '{' ? is_obj | (1ull << depth) : is_obj & ~(1ull << depth);
            expect_key = c == '{';
            want = W_NONE; key_ok = false;
        } else if (c == '}' || c == ']') {
            if (depth == 0) bad = true; else depth--;
            expect_key = false; want = W_NONE; key_ok = false;
        } else if (c == ',') {
            expect_key = depth < 64 && ((is_obj >> depth) & 1ull);
            want = W_NONE; key_ok = false;
        } else {
            if (!opened) bad = true;          // a line must start with '{'
            want = W_NONE;                    // a literal or a number: not a string value
        }
    }
    return opened && depth == 0 && !in_str && !bad;
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
        if (!json_scan_line(t, p, end, f)) bad = kJsonMalformed;
        if (!bad && (f.type.escaped || f.uid.escaped || f.ns.escaped || f.rv.escaped || f.phase.escaped)) bad = kJsonEscapedField;
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

// A window of the text held in shared memory, addressed by offsets in the text.
struct SmemText {
    const char* base; u64 off;
    __device__ __forceinline__ char operator[](u64 i) const { return base[i - off]; }
};

// Lines are a few hundred bytes of text each and a thread scans its line byte by byte: straight from global
// memory that is one sector per lane and step.  So a warp first copies the text of its next lines (they are
// consecutive in the log, hence one contiguous range) into shared memory with coalesced 16-byte loads, and every
// lane then scans its line there.  A line longer than the warp's window is scanned in place.
static constexpr int kJsonWarps = 4, kJsonWindow = 27 * 1024;      // bytes of text per warp and round (two blocks per SM)

__global__ void __launch_bounds__(kJsonWarps * 32) json_parse_kernel(JsonArgs a, u32* __restrict__ ticket) {
    extern __shared__ __align__(16) char s_text[];
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    char* win = s_text + warp * (kJsonWindow + 16);
    for (;;) {
        u64 g0 = 0;
        if (lane == 0) g0 = (u64)atomicAdd(ticket, 1u) * 32;
        g0 = __shfl_sync(0xffffffffu, g0, 0);
        if (g0 >= a.n_lines) break;
        const u64 g1 = min(a.n_lines, g0 + 32);
        u64 s = g0;
        while (s < g1) {
            // as many of the remaining lines of the group as fit into the window
            const u64 beg = s ? a.nl[s - 1] + 1 : 0;
            const u64 li = s + lane;
            const u64 my_end = li < g1 ? (li < a.n_nl ? a.nl[li] : a.len) : 0;
            const u32 fits = __ballot_sync(0xffffffffu, li < g1 && my_end - beg <= (u64)kJsonWindow);
            const u32 c = fits == 0xFFFFFFFFu ? 32u : (u32)__ffs(~fits) - 1;     // lines s .. s+c-1 fit (fits is a prefix of ones)
            if (c == 0) {
                if (lane == 0) json_line(a, a.text, s, beg, my_end);      // one very long line: scanned in place
                s += 1;
                __syncwarp();
                continue;
            }
            const u64 end_all = __shfl_sync(0xffffffffu, my_end, c - 1);
            const u64 abeg = beg & ~15ull;               // the text buffer is 256-byte aligned and padded
            for (u64 o = abeg + lane * 16; o < end_all; o += 512)
                *reinterpret_cast<uint4*>(win + (o - abeg)) = __ldcs(reinterpret_cast<const uint4*>(a.text + o));
            __syncwarp();
            if (lane < c) {
                const u64 p = li ? a.nl[li - 1] + 1 : 0;
                json_line(a, SmemText{win, abeg}, li, p, my_end);
            }
            __syncwarp();
            s += c;
        }
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
