// The watch-JSON line scanner: where in a line are the five strings the path reads
// (event['type'], metadata.uid / .namespace / .resourceVersion, status.phase: watcher/pod_watcher.py:162-167,
// :216-217, :267) and is pod.status truthy (:167, :209).
//
// Plain C++ on purpose, no CUDA in it: the device kernel (pte_json.cuh) and the CPU test harness
// (tests/json_scan_harness.cpp, checked against Python's json module) compile the very same function.
//
// One thread scans one line, 32 bytes at a time, in registers: byte classes become 32-bit masks by SWAR
// arithmetic that is the same for every byte and every thread (no divergence), the in-string mask is a prefix
// XOR over the unescaped quotes, and only brackets and colons OUTSIDE strings are then visited one by one --
// and not even those while the nesting is too deep to matter (depth > 3: labels, annotations, containers,
// conditions).  The byte-at-a-time state machine this replaces executed 5 400 warp instructions per line with 7
// of 32 lanes active (profiles/r02_json/): every lane sat in a different branch of it.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define PTE_JS_HD __host__ __device__ __forceinline__
#else
#define PTE_JS_HD inline
#endif
#if defined(__CUDA_ARCH__)
#define PTE_JS_UNROLL _Pragma("unroll")
#else
#define PTE_JS_UNROLL
#endif

namespace ptejson {
typedef unsigned long long u64;
typedef unsigned int u32;

struct Span { u64 s; u32 len; bool present; };
struct Fields { Span type, uid, ns, rv, phase; bool has_status; };

PTE_JS_HD u32 ctz32(u32 x) {
#if defined(__CUDA_ARCH__)
    return (u32)__ffs((int)x) - 1u;
#else
    return (u32)__builtin_ctz(x);
#endif
}
PTE_JS_HD u32 top32(u32 x) {      // index of the highest set bit, x != 0
#if defined(__CUDA_ARCH__)
    return 31u - (u32)__clz((int)x);
#else
    return 31u - (u32)__builtin_clz(x);
#endif
}
PTE_JS_HD u32 popc32(u32 x) {
#if defined(__CUDA_ARCH__)
    return (u32)__popc(x);
#else
    return (u32)__builtin_popcount(x);
#endif
}
PTE_JS_HD bool is_ws(char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\n'; }

// four bytes of w equal to the (7-bit) byte replicated in pat -> 4-bit mask, bit k for byte k.
// (w ^ pat) & 0x7f.. is zero in a matching byte, adding 0x7f leaves bit 7 clear only there; bit 7 of w itself must
// be clear too.  The multiply gathers bits 7, 15, 23, 31 into the top nibble (all partial products land on
// distinct bits, so nothing carries).
PTE_JS_HD u32 eq4(u32 w, u32 pat) {
    const u32 a = ((w ^ pat) & 0x7F7F7F7Fu) + 0x7F7F7F7Fu;
    const u32 z = ~(a | w) & 0x80808080u;
    return (z * 0x00204081u) >> 28;
}
PTE_JS_HD u32 eq32(const u32 (&w)[8], u32 pat) {
    u32 m = 0;
PTE_JS_UNROLL
    for (int k = 0; k < 8; k++) m |= eq4(w[k], pat) << (4 * k);
    return m;
}

enum : u32 { C_OTHER = 0, C_ROOT = 1, C_OBJECT = 2, C_META = 3, C_STATUS = 4 };
enum : u32 { W_NONE = 0, W_TYPE = 1, W_UID = 2, W_NS = 3, W_RV = 4, W_PHASE = 5, W_OBJECT = 6, W_META = 7, W_STATUS = 8 };

template <typename T>
PTE_JS_HD bool bytes_are(const T& t, u64 s, const char* name, u32 n) {
    for (u32 i = 0; i < n; i++) if (t[s + i] != name[i]) return false;
    return true;
}

// the key [s, s+len) seen in a container of meaning ctx: which value follows?  A key spelled with escapes never
// matches (its bytes differ from the plain name), as in the reference's client, which compares decoded names --
// nobody escapes these names, and one who did would simply not be seen.
template <typename T>
PTE_JS_HD u32 key_meaning(const T& t, u64 s, u32 len, u32 ctx) {
    if (ctx == C_ROOT) {
        if (len == 4 && bytes_are(t, s, "type", 4)) return W_TYPE;
        if (len == 6 && bytes_are(t, s, "object", 6)) return W_OBJECT;
    } else if (ctx == C_OBJECT) {
        if (len == 8 && bytes_are(t, s, "metadata", 8)) return W_META;
        if (len == 6 && bytes_are(t, s, "status", 6)) return W_STATUS;
    } else if (ctx == C_META) {
        if (len == 3 && bytes_are(t, s, "uid", 3)) return W_UID;
        if (len == 9 && bytes_are(t, s, "namespace", 9)) return W_NS;
        if (len == 15 && bytes_are(t, s, "resourceVersion", 15)) return W_RV;
    } else if (ctx == C_STATUS) {
        if (len == 5 && bytes_are(t, s, "phase", 5)) return W_PHASE;
    }
    return W_NONE;
}

PTE_JS_HD void set_field(Fields& f, u32 w, const Span& sp) {
    if (w == W_TYPE) f.type = sp; else if (w == W_UID) f.uid = sp; else if (w == W_NS) f.ns = sp;
    else if (w == W_RV) f.rv = sp; else if (w == W_PHASE) f.phase = sp;
}

// Line [p, end) of the text.  T gives bytes by text offset (t[i]) and 32-byte blocks by aligned offset
// (t.load32(blk, w): eight little-endian words; the text buffer is padded so that the block holding end-1 is readable).
// Returns false for a line that is not one JSON container: unbalanced brackets, an unterminated string, anything
// but '{' or '[' first.  Inside, the scan is as lenient as a scanner that skips what it does not need can be: it never
// looks at numbers, literals or commas.
// Semantics of the fields: a key counts where the reference would read it -- root.type, root.object.metadata.{uid,
// namespace,resourceVersion}, root.object.status.phase, each container reached through exactly that key; a value that
// is not a string leaves the field absent; of duplicate keys the last one wins (as in json.loads).
template <typename T>
PTE_JS_HD bool scan_line(const T& t, u64 p, u64 end, Fields& f) {
    const Span none = {0, 0, false};
    f.type = f.uid = f.ns = f.rv = f.phase = none;
    f.has_status = false;
    while (p < end && is_ws(t[p])) p++;
    if (p >= end) return false;
    { const char c = t[p]; if (c != '{' && c != '[') return false; }

    const u64 kNever = ~0ull;
    u32 depth = 0, ctxs = 0, want = W_NONE, pend = W_NONE;
    u64 want_pos = 0, pend_q = 0, last_open_q = kNever, last_close_q = kNever;
    bool in_str = false, esc_carry = false, bad = false;

    u32 w[8];
    u64 blk = p & ~31ull;
    t.load32(blk, w);
    for (; blk < end; blk += 32) {
        u32 nx[8];
        const bool more = blk + 32 < end;
        if (more) t.load32(blk + 32, nx);            // in flight while this block is looked at
        const u32 lo = blk < p ? (u32)(p - blk) : 0u;
        const u64 left = end - blk;
        u32 valid = left >= 32 ? 0xFFFFFFFFu : ((1u << (u32)left) - 1u);
        valid &= ~((1u << lo) - 1u);

        u32 w20[8];
PTE_JS_UNROLL
        for (int k = 0; k < 8; k++) w20[k] = w[k] | 0x20202020u;
        u32 Q = eq32(w, 0x22222222u) & valid;                 // "
        const u32 B = eq32(w, 0x5C5C5C5Cu) & valid;           // backslash
        const u32 O = eq32(w20, 0x7B7B7B7Bu) & valid;         // { [
        const u32 C = eq32(w20, 0x7D7D7D7Du) & valid;         // } ]
        const u32 K = eq32(w, 0x3A3A3A3Au) & valid;           // :

        // which bytes are escaped: every backslash that is not itself escaped escapes its successor
        if (B | (u32)esc_carry) {
            u32 escaped = esc_carry ? (1u << lo) : 0u;         // (lo: the first valid byte; a carry only exists past the first block)
            u32 m = B & ~escaped;
            esc_carry = false;
            while (m) {
                const u32 i = ctz32(m);
                if (i == 31) { esc_carry = true; m = 0; }
                else { escaped |= 2u << i; m &= ~(3u << i); }
            }
            Q &= ~escaped;
        }
        // s: bit i set = byte i is inside a string (opening quote included, closing quote not)
        u32 s = Q;
        s ^= s << 1; s ^= s << 2; s ^= s << 4; s ^= s << 8; s ^= s << 16;
        if (in_str) s = ~s;
        in_str = (s >> 31) != 0;
        const u32 openq = Q & s, closeq = Q & ~s;

        // a field's value string opened earlier (at pend_q) and not closed yet: the next quote behind pend_q closes it
        if (pend != W_NONE && Q && pend_q < blk + 32) {
            u32 m = Q;
            if (pend_q >= blk) m &= ~((2u << (u32)(pend_q - blk)) - 1u);
            if (m) {
                const u64 e = blk + ctz32(m);
                const Span sp = {pend_q + 1, (u32)(e - pend_q - 1), true};
                set_field(f, pend, sp);
                pend = W_NONE;
            }
        }

        const u32 Os = O & ~s, Cs = C & ~s;
        const u32 n_close = popc32(Cs);
        if (depth >= 4 + n_close) {
            // too deep for anything in this block to matter: only the nesting changes
            depth += popc32(Os) - n_close;
        } else {
            u32 ev = Os | Cs | (K & ~s);
            while (ev) {
                const u32 i = ctz32(ev);
                const u32 bit = 1u << i;
                ev &= ev - 1;
                const u64 pos = blk + i;
                if (Os & bit) {
                    if (depth < 3) {
                        u32 nc = C_OTHER;
                        if (t[pos] == '{') {
                            if (depth == 0) nc = C_ROOT;
                            else if (want != W_NONE) {
                                bool adjacent = true;               // the container must BE the value of that key
                                for (u64 q = want_pos + 1; q < pos; q++) adjacent = adjacent && is_ws(t[q]);
                                if (adjacent) {
                                    if (want == W_OBJECT) nc = C_OBJECT;
                                    else if (want == W_META) nc = C_META;
                                    else if (want == W_STATUS) { nc = C_STATUS; f.has_status = true; }   // pod.status is truthy
                                }
                            }
                        }
                        ctxs = (ctxs & ~(15u << (4 * (depth + 1)))) | (nc << (4 * (depth + 1)));
                    }
                    depth++;
                    want = W_NONE;
                } else if (Cs & bit) {
                    if (depth == 0) bad = true; else depth--;
                    want = W_NONE;
                } else if (depth >= 1 && depth <= 3) {
                    want = W_NONE;
                    const u32 ctx = (ctxs >> (4 * depth)) & 15u;
                    if (ctx != C_OTHER) {
                        // the key: the last complete string in front of this colon
                        const u32 below = bit - 1u;
                        const u64 ks = (openq & below) ? blk + top32(openq & below) : last_open_q;
                        const u64 ke = (closeq & below) ? blk + top32(closeq & below) : last_close_q;
                        if (ks != kNever && ke != kNever && ke > ks) {
                            const u32 wv = key_meaning(t, ks + 1, (u32)(ke - ks - 1), ctx);
                            if (wv >= W_OBJECT) { want = wv; want_pos = pos; }
                            else if (wv != W_NONE) {
                                set_field(f, wv, none);              // a later duplicate wins, string or not
                                pend = W_NONE;
                                u64 q = pos + 1;
                                while (q < end && is_ws(t[q])) q++;
                                if (q < end && t[q] == '"') {
                                    pend = wv; pend_q = q;
                                    if (q < blk + 32) {
                                        const u32 m = Q & ~((2u << (u32)(q - blk)) - 1u);
                                        if (m) {
                                            const u64 e = blk + ctz32(m);
                                            const Span sp = {q + 1, (u32)(e - q - 1), true};
                                            set_field(f, wv, sp);
                                            pend = W_NONE;
                                        }
                                    }
                                }
                            }
                        }
                    }
                } else {
                    want = W_NONE;
                }
            }
        }
        if (openq) last_open_q = blk + top32(openq);
        if (closeq) last_close_q = blk + top32(closeq);
        if (more) {
PTE_JS_UNROLL
            for (int k = 0; k < 8; k++) w[k] = nx[k];
        }
    }
    return depth == 0 && !in_str && !bad && pend == W_NONE;
}

}  // namespace ptejson
