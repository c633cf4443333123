// CPU build of the device's JSON line scanner (k8s_watcher_b200/csrc/pte_json_scan.h), for tests only:
// tests/test_json_scan_cpu.py compiles this with g++ and checks the scanner against Python's json module.
#include <string.h>
#include "../k8s_watcher_b200/csrc/pte_json_scan.h"

struct HostText {
    const char* base;
    char operator[](unsigned int i) const { return base[i]; }
    void load32(unsigned int blk, unsigned int (&w)[8]) const { memcpy(w, base + blk, 32); }
    unsigned long long load8(unsigned int s) const { unsigned long long v; memcpy(&v, base + s, 8); return v; }
};

// out: for each of type, uid, ns, rv, phase: start, length (or -1, -1 when absent); then has_status.  Returns 1 when the line is accepted.
extern "C" int scan_line_host(const char* text, unsigned long long p, unsigned long long end, long long* out) {
    ptejson::Fields f;
    struct HostQueue {
        unsigned int w[ptejson::kQueueCap][4];
        void put(unsigned int e, unsigned int k, unsigned int v) { w[e][k] = v; }
        unsigned int get(unsigned int e, unsigned int k) const { return w[e][k]; }
    } q;
    const bool ok = ptejson::scan_line(HostText{text}, (unsigned int)p, (unsigned int)end, f, q);
    const ptejson::Span sp[5] = {f.type, f.uid, f.ns, f.rv, f.phase};
    for (int k = 0; k < 5; k++) {
        out[2 * k] = ptejson::span_present(sp[k]) ? (long long)ptejson::span_start(sp[k]) : -1;
        out[2 * k + 1] = ptejson::span_present(sp[k]) ? (long long)ptejson::span_len(sp[k]) : -1;
    }
    out[10] = f.has_status ? 1 : 0;
    return ok ? 1 : 0;
}

// the bytes the string body [s, s + raw_len) stands for -> out (cap bytes); returns their number, -1 for a broken escape
extern "C" long long unescape_host(const char* text, unsigned long long s, unsigned long long raw_len, unsigned char* out, unsigned long long cap) {
    const HostText t{text};
    ptejson::Unescape<HostText> u(t, (unsigned int)s);
    unsigned long long n = 0;
    while (u.pos < s + raw_len || u.n_queue) {
        const unsigned int b = u.next();
        if (u.bad) return -1;
        if (n < cap) out[n] = (unsigned char)b;
        n++;
    }
    return u.pos > s + raw_len ? -1 : (long long)n;
}
