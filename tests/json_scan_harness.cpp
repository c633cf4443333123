// CPU build of the device's JSON line scanner (k8s_watcher_b200/csrc/pte_json_scan.h), for tests only:
// tests/test_json_scan_cpu.py compiles this with g++ and checks the scanner against Python's json module.
#include <string.h>
#include "../k8s_watcher_b200/csrc/pte_json_scan.h"

struct HostText {
    const char* base;
    char operator[](unsigned long long i) const { return base[i]; }
    void load32(unsigned long long blk, unsigned int (&w)[8]) const { memcpy(w, base + blk, 32); }
};

// out: for each of type, uid, ns, rv, phase: start, length (or -1, -1 when absent); then has_status.  Returns 1 when the line is accepted.
extern "C" int scan_line_host(const char* text, unsigned long long p, unsigned long long end, long long* out) {
    ptejson::Fields f;
    const bool ok = ptejson::scan_line(HostText{text}, p, end, f);
    const ptejson::Span* sp[5] = {&f.type, &f.uid, &f.ns, &f.rv, &f.phase};
    for (int k = 0; k < 5; k++) {
        out[2 * k] = sp[k]->present ? (long long)sp[k]->s : -1;
        out[2 * k + 1] = sp[k]->present ? (long long)sp[k]->len : -1;
    }
    out[10] = f.has_status ? 1 : 0;
    return ok ? 1 : 0;
}
