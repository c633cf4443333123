/*
 * oracle.c -- CPU restatement of the bulk phase-transition path, for scale.
 * TEST INFRASTRUCTURE ONLY: nothing under k8s_watcher_b200/ links or loads this;
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs do.
 *
 * What it restates (sequentially, one event at a time, the way the reference
 * watcher's loop at /root/reference/watcher/pod_watcher.py:266-269 runs):
 *   - should_process_event   pod_watcher.py:204-212   -> survives()
 *   - namespace allow-list   pod_watcher.py:226-229   -> survives()
 *   - "Unknown" substitution pod_watcher.py:167       -> eff_phase()
 *   - last-known-phase layer: ABSENT upstream; SURVEY.md section 8a row 7
 *     (PARITY UNPINNED for this layer; pinned against oracle/podwatch_oracle.py,
 *     whose filter half is pinned to the real reference by tests/golden).
 *
 * It also holds an independent, strictly sequential implementation of the
 * synthetic replay-log generator (spec in include/pte.h, "Synthetic replay
 * logs"), so that the engine's device generator can be checked against it.
 *
 * Build: make -C oracle   (gcc -O2 -shared -fPIC -> oracle/liboracle.so)
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef uint64_t u64;
typedef uint32_t u32;
typedef uint8_t u8;

enum { T_ADDED = 0, T_MODIFIED = 1, T_DELETED = 2 };
enum { PH_SUCCEEDED = 2, PH_FAILED = 3, PH_UNKNOWN = 4, PH_NULL = 5, PH_NOSTATUS = 6 };

static inline u64 mix64(u64 z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static inline u64 mulhi64(u64 a, u64 b) { return (u64)(((unsigned __int128)a * b) >> 64); }

#define UID_A 0x6B38735F77617463ull
#define UID_B 0x5F62323030706F64ull

/* ---------------- synthetic log: one event, counter-based part ------------- */
typedef struct {
    u64 seed, n_pods_local;
    u32 dist, p_del_q24, p_nostatus_q24, p_null_q24, shard, n_shards;
} gen_params;

typedef struct { u64 q, pod, rv, uid_hi, uid_lo; u32 ns; u8 phase, draw_del; } gen_event;

static inline u32 bitlen64(u64 x) { u32 n = 0; while (x) { n++; x >>= 1; } return n; }

static inline void gen_one(const gen_params* g, u64 j, gen_event* o) {
    u64 e = j * g->n_shards + g->shard;
    u64 r1 = mix64(g->seed ^ mix64(e + 0x9E3779B97F4A7C15ull));
    u64 r2 = mix64(r1 + 0xD1B54A32D192ED03ull);
    u64 r3 = mix64(r2 + 0x8CB92BA72F3D8DD7ull);
    u64 r4 = mix64(r3 + 0x2545F4914F6CDD1Dull);
    u64 P = g->n_pods_local, q;
    if (g->dist == 0) {
        q = mulhi64(r1, P);
    } else {
        u32 nb = bitlen64(P);
        u64 oct = mulhi64(r1, nb);
        u64 width = 1ull << oct;
        u64 rank = (width - 1) + (r2 & (width - 1));
        if (rank >= P) rank %= P;
        q = (rank * 2654435761ull + 40503ull) % P;
    }
    o->q = q;
    o->pod = q * g->n_shards + g->shard;
    o->rv = e + 1;
    u8 ph = (u8)(((r3 & 0xFFFFFFFFull) * 5) >> 32);
    if ((u32)(r4 & 0xFFFFFF) < g->p_nostatus_q24) ph = PH_NOSTATUS;
    else if ((u32)((r4 >> 24) & 0xFFFFFF) < g->p_null_q24) ph = PH_NULL;
    o->phase = ph;
    o->draw_del = (u32)(r3 >> 40) < g->p_del_q24;
    o->uid_hi = mix64(o->pod ^ UID_A);
    o->uid_lo = mix64(~o->pod ^ UID_B);
    o->ns = (u32)(o->pod % 5);
}

/* type rule (stateful, per pod): first event of a pod, or the event after a
 * delete-draw event of that pod, is ADDED; otherwise a delete-draw event is
 * DELETED; otherwise MODIFIED.  st: 0 never seen, 1 previous had no draw, 2 previous had draw. */
static inline u8 gen_type(u8* st, u64 q, u8 draw) {
    u8 s = st[q], t;
    if (s != 1) t = T_ADDED; else t = draw ? T_DELETED : T_MODIFIED;
    st[q] = draw ? 2 : 1;
    return t;
}

void orc_generate(u64 seed, u64 n, u64 n_pods_local, u32 dist, u32 p_del_q24, u32 p_nostatus_q24,
                  u32 p_null_q24, u32 shard, u32 n_shards,
                  u64* uid_hi, u64* uid_lo, u64* rv, u32* ns, u8* phase, u8* type) {
    gen_params g = { seed, n_pods_local, dist, p_del_q24, p_nostatus_q24, p_null_q24, shard, n_shards };
    u8* st = (u8*)calloc(n_pods_local ? n_pods_local : 1, 1);
    gen_event ev;
    for (u64 j = 0; j < n; j++) {
        gen_one(&g, j, &ev);
        uid_hi[j] = ev.uid_hi; uid_lo[j] = ev.uid_lo; rv[j] = ev.rv; ns[j] = ev.ns; phase[j] = ev.phase;
        type[j] = gen_type(st, ev.q, ev.draw_del);
    }
    free(st);
}

/* shard function: inverse of uid_hi = mix64(pod ^ UID_A), then mod n_shards */
static inline u64 unxorshift(u64 x, int s) { u64 r = x; for (int k = s; k < 64; k += s) r = x ^ (r >> s); return r; }
u64 orc_unmix64(u64 z) {
    z = unxorshift(z, 31); z *= 0x319642B2D24D8EC3ull;
    z = unxorshift(z, 27); z *= 0x96DE1B173F119089ull;
    z = unxorshift(z, 30);
    return z;
}
u32 orc_shard_of(u64 uid_hi, u32 n_shards) { return (u32)((orc_unmix64(uid_hi) ^ UID_A) % n_shards); }

/* ---------------- filters ---------------- */
typedef struct { int critical_only; const u32* allow; int n_allow; } filt;

static inline int survives(const filt* f, u8 type, u8 phase, u32 ns) {
    /* pod_watcher.py:209: drop iff critical_only && type != DELETED && status truthy && phase not in [Failed, Succeeded] */
    if (f->critical_only && type != T_DELETED && phase != PH_NOSTATUS && phase != PH_FAILED && phase != PH_SUCCEEDED)
        return 0;
    /* pod_watcher.py:227: drop iff allow-list non-empty and namespace not in it */
    if (f->n_allow > 0) {
        int ok = 0;
        for (int k = 0; k < f->n_allow; k++) if (f->allow[k] == ns) { ok = 1; break; }
        if (!ok) return 0;
    }
    return 1;
}
static inline u8 eff_phase(u8 ph) { return ph == PH_NOSTATUS ? PH_UNKNOWN : ph; }   /* pod_watcher.py:167 */

/* ---------------- last-known-phase table: uid128 -> state byte ---------------- */
typedef struct { u64* hi; u64* lo; u8* val; u64 mask, used; } ptab;   /* val: 0 empty slot, 1 absent(deleted), 2+eff present */

static void ptab_init(ptab* t, u64 cap_hint) {
    u64 cap = 1024; while (cap < cap_hint * 2) cap <<= 1;
    t->hi = (u64*)calloc(cap, 8); t->lo = (u64*)calloc(cap, 8); t->val = (u8*)calloc(cap, 1);
    t->mask = cap - 1; t->used = 0;
}
static void ptab_free(ptab* t) { free(t->hi); free(t->lo); free(t->val); }
static void ptab_grow(ptab* t);
static inline u8* ptab_slot(ptab* t, u64 hi, u64 lo) {
    if (t->used * 2 > t->mask) ptab_grow(t);
    u64 i = mix64(hi ^ mix64(lo)) & t->mask;
    for (;;) {
        if (t->val[i] == 0) { t->hi[i] = hi; t->lo[i] = lo; t->val[i] = 1; t->used++; return &t->val[i]; }
        if (t->hi[i] == hi && t->lo[i] == lo) return &t->val[i];
        i = (i + 1) & t->mask;
    }
}
static void ptab_grow(ptab* t) {
    ptab n; ptab_init(&n, (t->mask + 1) * 2);
    for (u64 i = 0; i <= t->mask; i++) if (t->val[i]) *ptab_slot(&n, t->hi[i], t->lo[i]) = t->val[i];
    ptab_free(t); *t = n;
}

/* one sequential step of SURVEY.md 8a row 7; returns 1 if the event is emitted */
static inline int step(ptab* t, u64 hi, u64 lo, u8 type, u8 phase, int dedupe) {
    if (!dedupe) return 1;
    u8* s = ptab_slot(t, hi, lo);
    if (type == T_DELETED) { *s = 1; return 1; }
    u8 want = (u8)(2 + eff_phase(phase));
    if (*s != want) { *s = want; return 1; }
    return 0;
}

static inline u64 rec_digest(u64 pos, u64 rv, u64 hi, u64 lo, u32 ns, u8 phase, u8 type) {
    u64 small = ((u64)ns << 16) | ((u64)phase << 8) | type;
    u64 h = mix64(lo ^ small);
    h = mix64(hi ^ h);
    h = mix64(rv ^ h);
    return mix64(((pos + 1) * 0x9E3779B97F4A7C15ull) ^ h);
}

typedef struct { u64 rv, i; } rvi;
static int cmp_rvi(const void* a, const void* b) {
    const rvi* x = (const rvi*)a; const rvi* y = (const rvi*)b;
    if (x->rv != y->rv) return x->rv < y->rv ? -1 : 1;
    return x->i < y->i ? -1 : (x->i > y->i);
}

/* Replay a column log.  out_idx (capacity n, may be NULL) receives emitted arrival
 * indices in replay order; stats = {n_in, n_survived, n_emitted, n_tombstones}; returns digest. */
u64 orc_replay(const u64* uid_hi, const u64* uid_lo, const u64* rv, const u32* ns, const u8* phase,
               const u8* type, u64 n, int critical_only, const u32* allow, int n_allow, int dedupe,
               u64 pods_hint, u64* out_idx, u64* stats) {
    filt f = { critical_only, allow, n_allow };
    ptab t; ptab_init(&t, pods_hint ? pods_hint : 1024);
    rvi* order = NULL;
    for (u64 i = 0; i + 1 < n; i++) if (rv[i] > rv[i + 1]) {
        order = (rvi*)malloc(n * sizeof(rvi));
        for (u64 k = 0; k < n; k++) { order[k].rv = rv[k]; order[k].i = k; }
        qsort(order, n, sizeof(rvi), cmp_rvi);
        break;
    }
    u64 surv = 0, emitted = 0, tomb = 0, dig = 0;
    for (u64 k = 0; k < n; k++) {
        u64 i = order ? order[k].i : k;
        if (!survives(&f, type[i], phase[i], ns[i])) continue;
        surv++;
        if (step(&t, uid_hi[i], uid_lo[i], type[i], phase[i], dedupe)) {
            dig += rec_digest(emitted, rv[i], uid_hi[i], uid_lo[i], ns[i], phase[i], type[i]);
            if (out_idx) out_idx[emitted] = i;
            emitted++; tomb += type[i] == T_DELETED;
        }
    }
    free(order); ptab_free(&t);
    stats[0] = n; stats[1] = surv; stats[2] = emitted; stats[3] = tomb;
    return dig;
}

/* Generate + replay + digest without materialising the log (1B-event parity). */
u64 orc_generate_replay(u64 seed, u64 n, u64 n_pods_local, u32 dist, u32 p_del_q24, u32 p_nostatus_q24,
                        u32 p_null_q24, u32 shard, u32 n_shards, int critical_only, const u32* allow,
                        int n_allow, int dedupe, u64* stats) {
    gen_params g = { seed, n_pods_local, dist, p_del_q24, p_nostatus_q24, p_null_q24, shard, n_shards };
    filt f = { critical_only, allow, n_allow };
    /* pods are dense here, so the last-phase table can be a flat array indexed by q */
    u8* gst = (u8*)calloc(n_pods_local ? n_pods_local : 1, 1);
    u8* last = (u8*)calloc(n_pods_local ? n_pods_local : 1, 1);   /* 0/1 absent, 2+eff present */
    u64 surv = 0, emitted = 0, tomb = 0, dig = 0;
    gen_event ev;
    for (u64 j = 0; j < n; j++) {
        gen_one(&g, j, &ev);
        u8 type = gen_type(gst, ev.q, ev.draw_del);
        if (!survives(&f, type, ev.phase, ev.ns)) continue;
        surv++;
        int emit = 1;
        if (dedupe) {
            if (type == T_DELETED) last[ev.q] = 1;
            else { u8 want = (u8)(2 + eff_phase(ev.phase)); if (last[ev.q] == want) emit = 0; else last[ev.q] = want; }
        }
        if (emit) {
            dig += rec_digest(emitted, ev.rv, ev.uid_hi, ev.uid_lo, ev.ns, ev.phase, type);
            emitted++; tomb += type == T_DELETED;
        }
    }
    free(gst); free(last);
    stats[0] = n; stats[1] = surv; stats[2] = emitted; stats[3] = tomb;
    return dig;
}
