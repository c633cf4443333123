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

// A string value: where its body starts in the text and how many RAW bytes it has, packed (start << 32 | len << 1 | 1)
// so that choosing which of the five fields to set costs five predicated moves; 0 = absent.  Texts are < 4 GiB.
typedef u64 Span;
PTE_JS_HD Span make_span(u32 s, u32 len) { return ((u64)s << 32) | ((u64)len << 1) | 1ull; }
PTE_JS_HD bool span_present(Span sp) { return (sp & 1ull) != 0; }
PTE_JS_HD u32 span_start(Span sp) { return (u32)(sp >> 32); }
PTE_JS_HD u32 span_len(Span sp) { return (u32)sp >> 1; }
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

// up to eight characters as the little-endian word a load8() of them gives
PTE_JS_HD constexpr u64 word_of(const char* s, int n) {
    u64 v = 0;
    for (int i = n - 1; i >= 0; i--) v = (v << 8) | (u64)(unsigned char)s[i];
    return v;
}
PTE_JS_HD constexpr u64 low_bytes(int n) { return n >= 8 ? ~0ull : ((1ull << (8 * n)) - 1ull); }

// the key [s, s+len) seen in a container of meaning ctx: which value follows?  Compared as one or two 8-byte words.
// A key spelled with escapes never matches (its bytes differ from the plain name): nobody escapes these names, and
// one who did would simply not be seen.
template <typename T>
PTE_JS_HD u32 key_meaning(const T& t, u32 s, u32 len, u32 ctx) {
    // the length alone rules most keys out (name, labels, annotations, spec, kind ...): no byte of those is looked at
    u32 w = W_NONE;
    u64 want = 0;
    if (ctx == C_ROOT) {
        if (len == 4) { w = W_TYPE; want = word_of("type", 4); }
        else if (len == 6) { w = W_OBJECT; want = word_of("object", 6); }
    } else if (ctx == C_OBJECT) {
        if (len == 8) { w = W_META; want = word_of("metadata", 8); }
        else if (len == 6) { w = W_STATUS; want = word_of("status", 6); }
    } else if (ctx == C_META) {
        if (len == 3) { w = W_UID; want = word_of("uid", 3); }
        else if (len == 9) { w = W_NS; want = word_of("namespac", 8); }
        else if (len == 15) { w = W_RV; want = word_of("resource", 8); }
    } else if (ctx == C_STATUS) {
        if (len == 5) { w = W_PHASE; want = word_of("phase", 5); }
    }
    if (w == W_NONE) return W_NONE;
    const u64 mask = len >= 8 ? ~0ull : ((1ull << (8 * len)) - 1ull);
    if ((t.load8(s) & mask) != want) return W_NONE;
    if (w == W_NS && t[s + 8] != 'e') return W_NONE;
    if (w == W_RV && (t.load8(s + 8) & low_bytes(7)) != word_of("Version", 7)) return W_NONE;
    return w;
}

PTE_JS_HD void set_field(Fields& f, u32 w, Span sp) {
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
// Candidate colons wait in a small per-thread queue (QS: q.put(entry, word, value) / q.get(entry, word); shared memory
// on the device) until the line is through, or the queue is full: looking at them right away means that in every block
// some lane of the warp drags the other 31 through the key comparison and the value look-up.  An entry is four words:
// where the key string starts, where the colon is, (key length | meaning of the container << 4 | distance from the colon
// to the next opening quote << 8, all ones = none seen), and where that string closes (kNever = not seen).
static constexpr u32 kQueueCap = 8, kNoDelta = 0xFFFFFFu;

template <typename T, typename QS>
PTE_JS_HD void flush_queue(const T& t, QS& q, u32& n_q, Fields& f) {
    for (u32 e = 0; e < n_q; e++) {
        const u32 ks1 = q.get(e, 0), pos = q.get(e, 1), meta = q.get(e, 2), q_close = q.get(e, 3);
        const u32 wv = key_meaning(t, ks1, meta & 15u, (meta >> 4) & 15u);
        if (wv == W_NONE || wv >= W_OBJECT) continue;
        Span sp = 0;                                    // a later duplicate wins, string or not
        const u32 delta = meta >> 8;
        if (delta != kNoDelta && q_close != ~0u) {
            const u32 q_open = pos + delta;
            bool adjacent = true;                       // the string must BE the value: nothing but whitespace after the colon
            for (u32 x = pos + 1; x < q_open; x++) adjacent = adjacent && is_ws(t[x]);
            if (adjacent) sp = make_span(q_open + 1, q_close - q_open - 1);
        }
        set_field(f, wv, sp);
    }
    n_q = 0;
}

template <typename T, typename QS>
PTE_JS_HD bool scan_line(const T& t, u32 p, u32 end, Fields& f, QS& q) {
    f.type = f.uid = f.ns = f.rv = f.phase = 0;
    f.has_status = false;
    while (p < end && is_ws(t[p])) p++;
    if (p >= end) return false;
    { const char c = t[p]; if (c != '{' && c != '[') return false; }

    const u32 kNever = ~0u;
    u32 depth = 0, ctxs = 0, n_q = 0;
    u32 pend = 0, pend_e = 0, pend_q = 0;            // pend: 0 nothing, 1 entry pend_e waits for its opening quote, 2 for the closing one (opened at pend_q)
    u32 last_open_q = kNever, last_close_q = kNever;
    bool in_str = false, esc_carry = false, bad = false;

    u32 w[8];
    u32 blk = p & ~31u;
    t.load32(blk, w);
    for (; blk < end; blk += 32) {
        u32 nx[8];
        const bool more = end - blk > 32;
        if (more) t.load32(blk + 32, nx);            // in flight while this block is looked at
        const u32 lo = blk < p ? p - blk : 0u;
        const u32 left = end - blk;
        u32 valid = left >= 32 ? 0xFFFFFFFFu : ((1u << left) - 1u);
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
            u32 escaped = esc_carry ? 1u : 0u;                 // a carry only arrives in a block that starts at its bit 0
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

        // the value of the last queued candidate: its opening quote, then its closing quote, wherever they turn up
        if (pend != 0 && Q) {
            u32 m = Q;
            if (pend == 1 && openq) {
                const u32 o = ctz32(openq);
                const u32 colon = q.get(pend_e, 1);
                const u32 delta = blk + o - colon;
                q.put(pend_e, 2, (q.get(pend_e, 2) & 0xFFu) | ((delta < kNoDelta ? delta : kNoDelta) << 8));
                pend = delta < kNoDelta ? 2u : 0u;
                pend_q = blk + o;
                m = o == 31 ? 0u : (Q & ~((2u << o) - 1u));
            }
            if (pend == 2 && m) { q.put(pend_e, 3, blk + ctz32(m)); pend = 0; }
        }

        // Brackets outside strings are walked one by one (they carry the nesting and say what a container means); colons
        // are not: the colons between two brackets all belong to the container open there, so they are collected by mask
        // and only those directly inside root / metadata / status are looked at, after the walk, when all lanes of the
        // warp are back together.
        const u32 Os = O & ~s, Cs = C & ~s, Ks = K & ~s;
        const u32 n_close = popc32(Cs);
        u32 cand_root = 0, cand_meta = 0, cand_status = 0;
        if (depth >= 4 + n_close) {
            // too deep for anything in this block to matter: only the nesting changes
            depth += popc32(Os) - n_close;
        } else {
            u32 br = Os | Cs, done = 0;
            for (;;) {
                const u32 i = br ? ctz32(br) : 32u;
                const u32 upto = i >= 32 ? 0xFFFFFFFFu : ((1u << i) - 1u);
                const u32 ctx = depth <= 3 ? (ctxs >> (4 * depth)) & 15u : (u32)C_OTHER;      // (nibble 0 is never set)
                const u32 seg = Ks & upto & ~done;
                if (ctx == C_ROOT) cand_root |= seg; else if (ctx == C_META) cand_meta |= seg; else if (ctx == C_STATUS) cand_status |= seg;
                if (i >= 32) break;
                const u32 bit = 1u << i;
                done = upto | bit;
                br &= br - 1;
                if (Cs & bit) {
                    if (depth == 0) bad = true; else depth--;
                } else {
                    if (depth < 3) {
                        const u32 pos = blk + i;
                        u32 nc = C_OTHER;
                        if (t[pos] == '{') {
                            if (depth == 0) nc = C_ROOT;
                            else if (ctx == C_ROOT || ctx == C_OBJECT) {
                                // the container means something only as the value of "object" / "metadata" / "status":
                                // key string, colon, container, with nothing but whitespace in between
                                const u32 below = bit - 1u;
                                const u32 ks = (openq & below) ? blk + top32(openq & below) : last_open_q;
                                const u32 ke = (closeq & below) ? blk + top32(closeq & below) : last_close_q;
                                if (ks != kNever && ke != kNever && ke > ks) {
                                    u32 colons = 0;
                                    bool clean = true;
                                    for (u32 q = ke + 1; q < pos; q++) { const char c = t[q]; colons += c == ':'; clean = clean && (c == ':' || is_ws(c)); }
                                    if (clean && colons == 1) {
                                        const u32 wv = key_meaning(t, ks + 1, ke - ks - 1, ctx);
                                        if (wv == W_OBJECT) nc = C_OBJECT;
                                        else if (wv == W_META) nc = C_META;
                                        else if (wv == W_STATUS) { nc = C_STATUS; f.has_status = true; }   // pod.status is truthy
                                    }
                                }
                            }
                        }
                        ctxs = (ctxs & ~(15u << (4 * (depth + 1)))) | (nc << (4 * (depth + 1)));
                    }
                    depth++;
                }
            }
        }
        u32 cand = cand_root | cand_meta | cand_status;
        while (cand) {
            const u32 i = ctz32(cand);
            const u32 bit = 1u << i;
            cand &= cand - 1;
            const u32 ctx = (cand_root & bit) ? (u32)C_ROOT : (cand_meta & bit) ? (u32)C_META : (u32)C_STATUS;
            // the key: the last complete string in front of this colon; only its length is looked at here
            const u32 below = bit - 1u;
            const u32 ks = (openq & below) ? blk + top32(openq & below) : last_open_q;
            const u32 ke = (closeq & below) ? blk + top32(closeq & below) : last_close_q;
            if (ks == kNever || ke == kNever || ke <= ks) continue;
            const u32 len = ke - ks - 1;
            const bool maybe = ctx == C_ROOT ? len == 4 : (ctx == C_META ? (len == 3 || len == 9 || len == 15) : len == 5);
            if (!maybe) continue;
            if (n_q == kQueueCap) flush_queue(t, q, n_q, f);
            // the string that follows (if it is the value is checked when the entry is looked at)
            u32 delta = kNoDelta, q_close = ~0u;
            pend = 1; pend_e = n_q;
            const u32 oq = i == 31 ? 0u : (openq & ~((2u << i) - 1u));
            if (oq) {
                const u32 o = ctz32(oq);
                delta = o - i;
                pend = 2; pend_q = blk + o;
                const u32 cq = o == 31 ? 0u : (Q & ~((2u << o) - 1u));
                if (cq) { q_close = blk + ctz32(cq); pend = 0; }
            }
            q.put(n_q, 0, ks + 1); q.put(n_q, 1, blk + i); q.put(n_q, 2, len | (ctx << 4) | (delta << 8)); q.put(n_q, 3, q_close);
            n_q++;
        }
        if (openq) last_open_q = blk + top32(openq);
        if (closeq) last_close_q = blk + top32(closeq);
        if (more) {
            PTE_JS_UNROLL
            for (int k = 0; k < 8; k++) w[k] = nx[k];
        }
    }
    flush_queue(t, q, n_q, f);
    return depth == 0 && !in_str && !bad;
}

// ---- escape sequences -------------------------------------------------------------------------------------
// The bytes a JSON string body stands for, one at a time: \" \\ \/ \b \f \n \r \t and \uXXXX (UTF-8 encoded, surrogate pairs
// joined).  `bad` is set by anything else behind a backslash and by a surrogate without its partner.
PTE_JS_HD int hex_digit(char c) {
    if (c >= '0' && c <= '9') return c - '0';
    if (c >= 'a' && c <= 'f') return c - 'a' + 10;
    if (c >= 'A' && c <= 'F') return c - 'A' + 10;
    return -1;
}
template <typename T>
struct Unescape {
    const T& t;
    u32 pos;          // next raw byte
    u32 queue;        // decoded bytes not handed out yet (low byte first)
    u32 n_queue;
    bool bad;
    PTE_JS_HD Unescape(const T& t_, u32 s) : t(t_), pos(s), queue(0), n_queue(0), bad(false) {}
    PTE_JS_HD int hex4(u32 at) const {
        int v = 0;
        for (u32 k = 0; k < 4; k++) { const int d = hex_digit(t[at + k]); if (d < 0) return -1; v = v * 16 + d; }
        return v;
    }
    // the next decoded byte; the caller stops by count or by watching pos reach the closing quote (with n_queue == 0)
    PTE_JS_HD u32 next() {
        if (n_queue) { const u32 b = queue & 0xFFu; queue >>= 8; n_queue--; return b; }
        const char c = t[pos++];
        if (c != '\\') return (u32)(unsigned char)c;
        const char e = t[pos++];
        switch (e) {
            case '"': return '"';
            case '\\': return '\\';
            case '/': return '/';
            case 'b': return 8;
            case 'f': return 12;
            case 'n': return 10;
            case 'r': return 13;
            case 't': return 9;
            case 'u': break;
            default: bad = true; return (u32)(unsigned char)e;
        }
        int cp = hex4(pos);
        if (cp < 0) { bad = true; return 0; }
        pos += 4;
        if (cp >= 0xDC00 && cp <= 0xDFFF) { bad = true; return 0; }
        if (cp >= 0xD800 && cp <= 0xDBFF) {
            const int lo = (t[pos] == '\\' && t[pos + 1] == 'u') ? hex4(pos + 2) : -1;
            if (lo < 0xDC00 || lo > 0xDFFF) { bad = true; return 0; }
            pos += 6;
            cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
        }
        if (cp < 0x80) return (u32)cp;
        if (cp < 0x800) { queue = 0x80u | (cp & 0x3F); n_queue = 1; return 0xC0u | (cp >> 6); }
        if (cp < 0x10000) { queue = (0x80u | ((cp >> 6) & 0x3F)) | ((0x80u | (cp & 0x3F)) << 8); n_queue = 2; return 0xE0u | (cp >> 12); }
        queue = (0x80u | ((cp >> 12) & 0x3F)) | ((0x80u | ((cp >> 6) & 0x3F)) << 8) | ((0x80u | (cp & 0x3F)) << 16);
        n_queue = 3;
        return 0xF0u | (cp >> 18);
    }
};

// decoded length of the string body [s, s + raw_len); sets bad like Unescape
template <typename T>
PTE_JS_HD u32 decoded_length(const T& t, u32 s, u32 raw_len, bool& bad) {
    Unescape<T> u(t, s);
    u32 n = 0;
    while (u.pos < s + raw_len || u.n_queue) { u.next(); n++; if (u.bad) break; }
    bad = u.bad || u.pos > s + raw_len;       // an escape cut off by the closing quote cannot happen (the quote would be escaped), but a count is cheap
    return n;
}

}  // namespace ptejson
