// Host side of libpte.so: device memory, kernel sequencing and the C ABI of include/pte.h.
// The path it runs in bulk is the reference's per-event loop, watcher/pod_watcher.py:266-269
// -> handle_pod_event (:214-241), plus the last-known-phase layer of SURVEY.md 8a row 7.
#include "../../include/pte.h"
#include "pte_kernels.cuh"
#include "pte_partition.cuh"
#include "pte_json.cuh"
#include "pte_tiny.cuh"

#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>

namespace {

thread_local std::string g_err;

constexpr u64 kPad = 8192;                       // every per-event buffer is padded to this many elements
constexpr u64 kMaxEvents = (1ull << 30) - 1;     // look-back words carry 30-bit counts

inline u64 round_up(u64 x, u64 m) { return (x + m - 1) / m * m; }
inline u32 bitlen_host(u64 x) { u32 n = 0; while (x) { n++; x >>= 1; } return n; }

// radix pass configurations (threads, records per thread); PTE_PASS_CFG picks one for tuning runs
struct PassCfg { int threads, items, ballot_items; };
constexpr PassCfg kPassCfgs[] = {{384, 16, 0}, {512, 12, 0}, {512, 12, 4}, {512, 12, 12}, {256, 16, 0}, {512, 16, 0}};
constexpr int kNumPassCfgs = sizeof(kPassCfgs) / sizeof(kPassCfgs[0]);

template <int T, int I, int B>
cudaError_t launch_pass(int grid, cudaStream_t st, const u64* in, u64* out, const u64* n_ptr, u64 n_fixed, int shift,
                        u32 mask, const u32* hist, u32* lookback, u32* ticket) {
    size_t smem = sizeof(RadixPassSmem<T, I>);
    radix_pass_kernel<T, I, B><<<grid, T, smem, st>>>(in, out, n_ptr, n_fixed, shift, mask, hist, lookback, ticket);
    return cudaGetLastError();
}
template <int T, int I, int B>
cudaError_t prepare_pass(int* occ) {
    size_t smem = sizeof(RadixPassSmem<T, I>);
    cudaError_t e = cudaFuncSetAttribute(radix_pass_kernel<T, I, B>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, radix_pass_kernel<T, I, B>, T, smem);
}

struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool load(std::string& err) {
        if (lib) return true;
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* nm : names) { lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL); if (lib) break; }
        if (!lib) { err = std::string("dlopen libnccl failed: ") + dlerror(); return false; }
        GetUniqueId = (decltype(GetUniqueId))dlsym(lib, "ncclGetUniqueId");
        CommInitRank = (decltype(CommInitRank))dlsym(lib, "ncclCommInitRank");
        AllGather = (decltype(AllGather))dlsym(lib, "ncclAllGather");
        CommDestroy = (decltype(CommDestroy))dlsym(lib, "ncclCommDestroy");
        GetErrorString = (decltype(GetErrorString))dlsym(lib, "ncclGetErrorString");
        if (!GetUniqueId || !CommInitRank || !AllGather || !CommDestroy || !GetErrorString) {
            err = "libnccl is missing an expected symbol"; return false;
        }
        return true;
    }
};
NcclApi g_nccl;

}  // namespace

struct pte_engine {
    int device = 0, num_sms = 0;
    u64 max_events = 0, max_pods = 0, cap_ev = 0;
    cudaStream_t stream = nullptr;
    // the log
    u64 *uid_hi = nullptr, *uid_lo = nullptr, *rv = nullptr;
    u32* ns = nullptr;
    u8 *phase = nullptr, *type = nullptr;
    u64 n_events = 0;
    // interning dictionary
    Slot* table = nullptr;
    Slot* table_spare = nullptr;                             // target of the next stream-dictionary rebuild (allocated on first use)
    u32 n_rebuilds = 0; float ms_last_rebuild = 0.f;
    u64 table_cap = 0;
    RecLayout L{};
    // sort
    u64* rec_pool = nullptr;                                 // 3 x cap_ev words: recA | recB | recC, or the 24-byte partition records
    u64 *recA = nullptr, *recB = nullptr, *recC = nullptr;   // interned records (round-major) + sort ping-pong
    // partition-first replay (pte_partition.cuh)
    u64 table_alloc = 0;                                     // slots allocated (>= table_cap, >= kParts * part_slice)
    u32 part_slice = 0;                                      // dictionary slots per partition
    u32 *part_hist = nullptr, *part_start = nullptr, *part_nsegs = nullptr;
    Segment* segs = nullptr;
    u8* seg_export = nullptr;
    u64 part_min_events = ~0ull;                             // auto mode: logs at least this long take the partition path (PTE_PART_MIN); off until it wins
    u32 seg_q8 = 304;                                        // segment length in 1/256 of the mean partition (PTE_SEG_Q8)
    int path_mode = 0;                                       // PTE_PATH: 0 auto, 1 always sort, 2 always partition
    bool retry_after_rebuild = false;
    bool force_sort = false;                                 // one run: the partitioned attempt overflowed something
    int last_path = 0;                                       // path of the last run (0 sort, 1 partition)
    int replay_occ = 1, scatter_smem = 0, replay_smem = 0;
    int scatter_cfg = 0, scatter_occ = 1, scatter_tile = 6144, scatter_threads = 512;
    u32 *bitmap = nullptr, *keepmap = nullptr;               // one emit / keep bit per arrival position
    u8 *state_old = nullptr, *state_new = nullptr;           // per dictionary slot: last known phase of the stream
    // small batches in one launch (pte_tiny.cuh): packed staging in pinned host memory and on the device
    bool force_stage_timing = false;                         // PTE_STAGE_TIMING=1
    bool tiny_on = true;                                     // PTE_TINY=0: every batch takes the bulk kernels
    u8 *tiny_in_host = nullptr, *tiny_in_dev = nullptr;      // uid_hi | uid_lo | rv | ns | phase | type of one batch
    u8 *tiny_out_host = nullptr, *tiny_out_dev = nullptr;    // Counters | out_idx | out_type of one batch
    bool tiny_staged = false;                                // the batch is in tiny_in_dev, not yet in the column arrays
    bool tiny_result = false;                                // tiny_out_host holds the notifications of the last run
    bool carry_valid = false;                                // dictionary + state hold a stream that can be continued
    LogColumns shadow{};                                     // rv-ordered copy of an unordered log (allocated on first use)
    u32* perm = nullptr;                                     // shadow row -> arrival index
    char *pay_buf = nullptr, *pay_blob = nullptr;            // serialised payloads; JSON fragments
    u64 pay_buf_cap = 0, pay_blob_cap = 0, pay_off_cap = 0, pay_tile_cap = 0;
    u32* pay_frag_off = nullptr; u64 pay_frag_cap = 0;
    u64 *pay_rec_off = nullptr, *pay_tile_off = nullptr;
    // raw watch-JSON ingestion
    char* json_text = nullptr; u64 json_text_cap = 0;
    u64* json_nl = nullptr; u64 json_nl_cap = 0;
    u32* json_chunk_cnt = nullptr; u64 json_chunk_cap = 0;
    u64* json_ns_keys = nullptr; u32* json_ns_off = nullptr; u64* json_ph_keys = nullptr; u32* json_ph_off = nullptr;
    u32 *json_err = nullptr, *json_ns_remap = nullptr; u8* json_ph_remap = nullptr;
    std::vector<std::string> json_names[3];                  // [0] namespaces by ns_id, [1] phase strings for codes >= 7, [2] non-UUID uid strings by reserved key - 1
    u64* json_uid_keys = nullptr; u32 *json_uid_off = nullptr, *json_uid_remap = nullptr; bool json_uid_dirty = true;
    // pte_json_stream: the logs ingested are the batches of ONE stream -- ids and codes of names persist, rows are numbered by arrival
    bool json_stream = false; u64 json_seq = 0;
    std::unordered_map<std::string, u32> json_ids[3];
    u64* d_minmax = nullptr;
    float ms_payload = 0.f; u64 payload_bytes = 0;            // device time of the last payload_write_kernel
    float ms_json[2] = {0.f, 0.f}; u64 json_bytes = 0;        // device times of the last pte_ingest_json
    bool in_reorder = false;
    u32 last_ordered = 1;
    u64 pods_total = 0, stream_last_rv = 0, batch_first_rv = 0, batch_last_rv = 0;
    u32 *lookback = nullptr, *hist = nullptr;
    u64 lookback_words = 0;
    int pass_cfg = 0, pass_occ = 1, pass_tile = 6144;
    u64 diff_round_bytes = 64ull << 20;      // bitmap bytes one diff round may touch
    bool full_cols_valid = false;            // out_rv / out_ns / out_phase hold the last run's records (filled on demand)
    bool key_valid = false;                  // out_key holds the keys of the last run
    bool no_invert = false;                  // PTE_NO_INVERT=1: always set emitted bits (tuning / tests)
    bool no_defer = false;                   // PTE_NO_DEFER=1: one diff launch (and one read of the records) per bitmap slice
    Counters* ctr = nullptr;
    // filters
    u32* ns_bitmap = nullptr;
    u32 ns_words = 0, ns_bitmap_cap = 0;
    std::vector<u32> ns_bitmap_host;
    int ns_filter_on = 0, critical_only = 0;
    // outputs
    u32 *out_idx = nullptr, *out_ns = nullptr;
    u64 *out_rv = nullptr, *out_key = nullptr;
    u8 *out_phase = nullptr, *out_type = nullptr;
    bool has_run = false;
    bool cols_valid = false;                                 // the output columns of the last run exist (not a PTE_RUN_BITMAP_ONLY run)
    cudaStream_t copy_stream = nullptr;                      // pte_replay_pipelined: host -> device copies run ahead of the replay
    bool ns_loaded = true;                                   // the ns column holds caller data (or generator output), not zeros
    Counters last{};
    // timing
    std::vector<cudaEvent_t> ev;
    u32 n_launches = 0, n_diff_rounds = 0;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pass_ev;
    size_t pass_ev_used = 0;
    bool stage_timing = true;                                // per-stage events: only for runs long enough for them not to matter (>= 2^20 events)
    // multi-GPU
    ncclComm_t comm = nullptr;
    int rank = 0, n_ranks = 1;
    u64 *d_counts_send = nullptr, *d_counts_recv = nullptr;
    std::string err;
};

#define FAIL(h, code, msg)                                  \
    do {                                                    \
        std::string m__ = (msg);                            \
        if (h) (h)->err = m__; else g_err = m__;            \
        return (code);                                      \
    } while (0)

#define CK(h, call)                                                                                         \
    do {                                                                                                    \
        cudaError_t e__ = (call);                                                                           \
        if (e__ != cudaSuccess)                                                                             \
            FAIL(h, PTE_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__));                      \
    } while (0)

namespace {

#define PTE_FOR_EACH_PASS_CFG(X) \
    X(0, 384, 16, 0) X(1, 512, 12, 0) X(2, 512, 12, 4) X(3, 512, 12, 12) X(4, 256, 16, 0) X(5, 512, 16, 0)

int launch_pass_cfg(pte_engine* h, const u64* in, u64* out, const u64* n_ptr, u64 n_fixed, int shift, u32 mask,
                    const u32* hist) {
    int grid = h->num_sms * h->pass_occ;
    u32* ticket = &h->ctr->ticket[1];
    cudaError_t e = cudaErrorInvalidValue;
    switch (h->pass_cfg) {
#define X(id, T, I, B) case id: e = launch_pass<T, I, B>(grid, h->stream, in, out, n_ptr, n_fixed, shift, mask, hist, h->lookback, ticket); break;
        PTE_FOR_EACH_PASS_CFG(X)
#undef X
    }
    if (e != cudaSuccess) FAIL(h, PTE_ERR_CUDA, std::string("radix_pass_kernel: ") + cudaGetErrorString(e));
    h->n_launches++;
    return PTE_OK;
}

int prepare_pass_cfg(pte_engine* h) {
    cudaError_t e = cudaErrorInvalidValue;
    switch (h->pass_cfg) {
#define X(id, T, I, B) case id: e = prepare_pass<T, I, B>(&h->pass_occ); break;
        PTE_FOR_EACH_PASS_CFG(X)
#undef X
    }
    if (e != cudaSuccess) FAIL(h, PTE_ERR_CUDA, std::string("radix pass set-up: ") + cudaGetErrorString(e));
    if (h->pass_occ < 1) FAIL(h, PTE_ERR_CUDA, "radix pass does not fit on an SM");
    h->pass_tile = kPassCfgs[h->pass_cfg].threads * kPassCfgs[h->pass_cfg].items;
    return PTE_OK;
}

// Stable LSD sort of the records in `src` on bits [bit_lo, bit_hi).  `src` is left intact: the first
// pass scatters into `x`, later passes ping-pong between `x` and `y`.
// n is read on the device from n_ptr when given (n_upper bounds it for sizing), else n_upper.
// On return *result points at the buffer that holds the sorted records.
int sort_field(pte_engine* h, u64* src, u64* x, u64* y, const u64* n_ptr, u64 n_upper, int bit_lo, int bit_hi,
               u64** result, bool hist_ready = false) {
    *result = src;
    if (bit_hi <= bit_lo || n_upper == 0) return PTE_OK;
    const int passes = (bit_hi - bit_lo + 7) / 8;
    if (passes > kMaxPasses) FAIL(h, PTE_ERR_INVALID, "sort field wider than 64 bits");
    if (!hist_ready) {     // the producer of the records may have counted the digits already
        CK(h, cudaMemsetAsync(h->hist, 0, sizeof(u32) * kMaxPasses * kRadix, h->stream));
        int hgrid = (int)std::min<u64>((u64)h->num_sms * 8, (n_upper + 511) / 512);
        radix_hist_kernel<256><<<std::max(hgrid, 1), 256, 0, h->stream>>>(src, n_ptr, n_upper, bit_lo, passes, bit_hi, h->hist);
        CK(h, cudaGetLastError());
        h->n_launches++;
    }
    const u64 tiles = (n_upper + h->pass_tile - 1) / h->pass_tile;
    if (tiles * kRadix > h->lookback_words) FAIL(h, PTE_ERR_CAPACITY, "look-back buffer too small");
    u64 *a = src, *b = x;
    for (int p = 0; p < passes; p++) {
        const int shift = bit_lo + 8 * p;
        const int rem = bit_hi - shift;
        const u32 mask = rem >= 8 ? 0xFFu : ((1u << rem) - 1);
        CK(h, cudaMemsetAsync(h->lookback, 0, tiles * kRadix * sizeof(u32), h->stream));
        CK(h, cudaMemsetAsync(&h->ctr->ticket[1], 0, sizeof(u32), h->stream));
        if (h->pass_ev_used == h->pass_ev.size()) {
            cudaEvent_t e0, e1;
            CK(h, cudaEventCreate(&e0));
            CK(h, cudaEventCreate(&e1));
            h->pass_ev.emplace_back(e0, e1);
        }
        auto& pe = h->pass_ev[h->pass_ev_used++];             // (counted in any case: pte_stats.n_passes)
        if (h->stage_timing) CK(h, cudaEventRecord(pe.first, h->stream));
        int rc = launch_pass_cfg(h, a, b, n_ptr, n_upper, shift, mask, h->hist + p * kRadix);
        if (rc) return rc;
        if (h->stage_timing) CK(h, cudaEventRecord(pe.second, h->stream));
        if (p == 0) { a = x; b = y; } else std::swap(a, b);
    }
    *result = a;
    return PTE_OK;
}

template <typename T>
int dev_alloc(pte_engine* h, T** p, u64 count) {
    CK(h, cudaMalloc((void**)p, count * sizeof(T)));
    return PTE_OK;
}

// (re)allocate a device buffer that only ever grows
template <typename T>
int grow(pte_engine* h, T** p, u64* cap, u64 need) {
    if (*cap >= need) return PTE_OK;
    if (*p) { CK(h, cudaFree(*p)); *p = nullptr; *cap = 0; }
    CK(h, cudaMalloc((void**)p, need * sizeof(T)));
    *cap = need;
    return PTE_OK;
}

int check_flags(pte_engine* h, u32 flags) {
    if (flags & kFlagBadEvent)
        FAIL(h, PTE_ERR_BAD_EVENT, "log holds an event with type not in {ADDED,MODIFIED,DELETED}, phase code > 62 or uid 0:0");
    if (flags & kFlagTableFull)
        FAIL(h, PTE_ERR_TABLE_FULL, "more distinct uids than the engine was created for (max_pods)");
    return PTE_OK;
}

// partition scatter configurations (threads, records per thread, CTAs per SM); PTE_SCATTER_CFG picks one for tuning runs
#define PTE_FOR_EACH_SCATTER_CFG(X) X(0, 512, 12, 1) X(1, 256, 12, 2) X(2, 256, 8, 3) X(3, 512, 8, 1)

template <int T, int I, int B>
cudaError_t prepare_scatter(pte_engine* h) {
    h->scatter_smem = (int)sizeof(ScatterSmem<T, I>);
    h->scatter_tile = T * I; h->scatter_threads = T;
    cudaError_t e = cudaFuncSetAttribute(part_scatter_kernel<T, I, B, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->scatter_smem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(part_scatter_kernel<T, I, B, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->scatter_smem);
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&h->scatter_occ, part_scatter_kernel<T, I, B, false>, T, h->scatter_smem);
    return e;
}
template <int T, int I, int B>
cudaError_t launch_scatter(pte_engine* h, int grid, bool filtered, const ScatterArgs& sa) {
    if (filtered) part_scatter_kernel<T, I, B, true><<<grid, T, h->scatter_smem, h->stream>>>(sa);
    else part_scatter_kernel<T, I, B, false><<<grid, T, h->scatter_smem, h->stream>>>(sa);
    return cudaGetLastError();
}

// Everything whose size follows max_events (log columns, record buffers, bitmaps, look-back words, output columns).
// Used by pte_create and by pte_grow; the dictionary and the per-pod state are not touched.
int alloc_event_buffers(pte_engine* h) {
    h->cap_ev = round_up(h->max_events, kPad) + kPad;
    const u64 c = h->cap_ev;
    int rc = 0;
    rc |= dev_alloc(h, &h->uid_hi, c); rc |= dev_alloc(h, &h->uid_lo, c); rc |= dev_alloc(h, &h->rv, c);
    rc |= dev_alloc(h, &h->ns, c); rc |= dev_alloc(h, &h->phase, c); rc |= dev_alloc(h, &h->type, c);
    rc |= dev_alloc(h, &h->rec_pool, 3 * c);
    rc |= dev_alloc(h, &h->bitmap, c / 32); rc |= dev_alloc(h, &h->keepmap, c / 32);
    // look-back words: one per (radix tile, digit); one per (partition-scatter tile, partition)
    const u64 min_tile = 2048;
    h->lookback_words = std::max<u64>((c + min_tile - 1) / min_tile * kRadix, (c + kInternTile - 1) / kInternTile);
    h->lookback_words = std::max<u64>(h->lookback_words, (c / 2048 + 1) * kParts);
    rc |= dev_alloc(h, &h->lookback, h->lookback_words);
    rc |= dev_alloc(h, &h->out_idx, c); rc |= dev_alloc(h, &h->out_ns, c);
    rc |= dev_alloc(h, &h->out_rv, c); rc |= dev_alloc(h, &h->out_key, c);
    rc |= dev_alloc(h, &h->out_phase, c); rc |= dev_alloc(h, &h->out_type, c);
    if (rc) return PTE_ERR_CUDA;
    h->recA = h->rec_pool; h->recB = h->rec_pool + c; h->recC = h->rec_pool + 2 * c;
    // pads must be readable garbage-free for the vector loads past n
    CK(h, cudaMemsetAsync(h->rv, 0, c * sizeof(u64), h->stream));
    CK(h, cudaMemsetAsync(h->ns, 0, c * sizeof(u32), h->stream));
    return PTE_OK;
}

int prepare_partition(pte_engine* h) {
    h->replay_smem = kRpStages * kRpBufRecs * 24 + 2 * (int)((h->part_slice + 15) & ~15u) + (int)sizeof(ReplaySmemFixed);
    cudaError_t e = cudaErrorInvalidValue;
    switch (h->scatter_cfg) {
#define X(id, T, I, B) case id: e = prepare_scatter<T, I, B>(h); break;
        PTE_FOR_EACH_SCATTER_CFG(X)
#undef X
    }
    if (e == cudaSuccess && h->scatter_occ < 1) e = cudaErrorInvalidConfiguration;
    if (e == cudaSuccess) e = cudaFuncSetAttribute(part_replay_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, h->replay_smem);
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&h->replay_occ, part_replay_kernel, kRpThreads, h->replay_smem);
    if (e != cudaSuccess) FAIL(h, PTE_ERR_CUDA, std::string("partition path set-up: ") + cudaGetErrorString(e));
    if (h->replay_occ < 1) h->path_mode = 1;      // the state bytes of one partition do not fit in shared memory: sort path only
    return PTE_OK;
}

// The T-trace of the resident log by the partition-first path: fills the emit bitmap (and keepmap when filtered).
// Events ev[6] (after the counts), ev[1] (after the scatter), ev[7] (before the replay) and ev[3] are recorded here.
int run_partitioned(pte_engine* h, const InternArgs& a, bool filtered) {
    cudaStream_t st = h->stream;
    const u64 n = h->n_events;
    const u32 NS = h->part_slice;
    if (!h->seg_export) CK(h, cudaMalloc((void**)&h->seg_export, (u64)kMaxSegments * NS));
    CK(h, cudaMemsetAsync(h->part_hist, 0, sizeof(u32) * kParts, st));
    if (filtered) {
        int fgrid = (int)std::min<u64>((n / 4 + 255) / 256 + 1, (u64)h->num_sms * 16);
        filter_flag_kernel<<<fgrid, 256, 0, st>>>(a, h->keepmap);
        CK(h, cudaGetLastError());
        h->n_launches++;
    }
    PartHistArgs ha{h->uid_lo, h->rv, h->keepmap, n, h->part_hist, h->ctr};
    int hgrid = (int)std::min<u64>((n / 4 + 255) / 256 + 1, (u64)h->num_sms * 8);
    if (filtered) part_hist_kernel<true><<<hgrid, 256, 0, st>>>(ha);
    else part_hist_kernel<false><<<hgrid, 256, 0, st>>>(ha);
    CK(h, cudaGetLastError());
    part_scan_kernel<<<1, kParts, 0, st>>>(h->part_hist, h->part_start, h->segs, h->part_nsegs, h->ctr, h->seg_q8, filtered ? 0 : 1);
    CK(h, cudaGetLastError());
    CK(h, cudaEventRecord(h->ev[6], st));
    const u64 tiles = (n + h->scatter_tile - 1) / h->scatter_tile;
    if (tiles * kParts > h->lookback_words) FAIL(h, PTE_ERR_CAPACITY, "look-back buffer too small");
    CK(h, cudaMemsetAsync(h->lookback, 0, tiles * kParts * sizeof(u32), st));
    ScatterArgs sa{h->uid_hi, h->uid_lo, h->phase, h->type, h->keepmap, n, h->part_start, h->rec_pool, h->lookback,
                   &h->ctr->ticket[2], h->ctr};
    int sgrid = (int)std::min<u64>(tiles, (u64)h->num_sms * h->scatter_occ);
    {
        cudaError_t e = cudaErrorInvalidValue;
        switch (h->scatter_cfg) {
#define X(id, T, I, B) case id: e = launch_scatter<T, I, B>(h, sgrid, filtered, sa); break;
            PTE_FOR_EACH_SCATTER_CFG(X)
#undef X
        }
        CK(h, e);
    }
    CK(h, cudaEventRecord(h->ev[1], st));
    CK(h, cudaEventRecord(h->ev[2], st));
    // polarity of the bitmap from a sample replay (a prefix of four partitions, no side effects but dictionary entries)
    ReplayArgs ra;
    ra.rec = h->rec_pool; ra.segs = h->segs + kMaxSegments; ra.n_segs = h->part_nsegs + 1; ra.ticket = &h->ctr->ticket[3];
    ra.first_seg = kMaxSegments; ra.table = h->table; ra.slice_slots = NS; ra.bitmap = h->bitmap; ra.ctr = h->ctr;
    ra.seg_export = h->seg_export; ra.pending = reinterpret_cast<PendingHead*>(h->out_key);
    ra.pending_cap = (u32)std::min<u64>(h->cap_ev / 2, 0xFFFFFFFFull); ra.count_only = 1;
    const bool sample = !h->no_invert;
    if (sample) {
        part_replay_kernel<<<kSampleSegments, kRpThreads, h->replay_smem, st>>>(ra);
        CK(h, cudaGetLastError());
        CK(h, cudaMemsetAsync(&h->ctr->ticket[3], 0, sizeof(u32), st));
        h->n_launches++;
    }
    bitmap_init_kernel<<<(int)std::min<u64>(((n + 31) / 32 + 255) / 256, (u64)h->num_sms * 8), 256, 0, st>>>(
        h->bitmap, filtered ? h->keepmap : nullptr, n, h->ctr, sample ? 1 : 0);
    CK(h, cudaGetLastError());
    CK(h, cudaEventRecord(h->ev[7], st));
    ra.segs = h->segs; ra.n_segs = h->part_nsegs; ra.first_seg = 0; ra.count_only = 0;
    part_replay_kernel<<<h->num_sms * h->replay_occ, kRpThreads, h->replay_smem, st>>>(ra);
    CK(h, cudaGetLastError());
#ifdef PTE_EXP_TRACE      // development build (-DPTE_EXP_TRACE): time stamps of one CTA's pipeline, tiles 200..223 of its first segment
    {
        static unsigned long long tr[64 * 8];
        cudaStreamSynchronize(st);
        cudaMemcpyFromSymbol(tr, g_rp_trace, sizeof(tr));
        fprintf(stderr, "tile: prober0 start-wait, tile landed, prober0 done, LAST prober done, replayer0 start, replayer0 done, LAST replayer done, refill issued (us, relative)\n");
        for (int t = 0; t < 24; t++) {
            const unsigned long long* r = tr + t * 8; const double b = (double)tr[0];
            fprintf(stderr, "%3d: %8.2f %8.2f %8.2f %8.2f %8.2f %8.2f %8.2f %8.2f\n", t, (r[0]-b)/1e3, (r[1]-b)/1e3, (r[2]-b)/1e3, (r[6]-b)/1e3, (r[3]-b)/1e3, (r[4]-b)/1e3, (r[7]-b)/1e3, (r[5]-b)/1e3);
            if (t == 23) { unsigned long long z[64 * 8] = {0}; cudaMemcpyToSymbol(g_rp_trace, z, sizeof(z)); }
        }
    }
#endif
    pending_fixup_kernel<<<h->num_sms * 4, 256, 0, st>>>(ra.pending, ra.pending_cap, h->segs, h->seg_export, NS, h->bitmap, h->ctr);
    CK(h, cudaGetLastError());
    h->n_launches += 6;
    return PTE_OK;
}

// Stream dictionary full (or nearly): keep the pods with a live phase, forget the rest.  Returns the live count.
int rebuild_stream_table(pte_engine* h) {
    cudaStream_t st = h->stream;
    if (!h->table_spare) CK(h, cudaMalloc((void**)&h->table_spare, h->table_alloc * sizeof(Slot)));
    cudaEvent_t e0, e1;
    CK(h, cudaEventCreate(&e0)); CK(h, cudaEventCreate(&e1));
    CK(h, cudaEventRecord(e0, st));
    CK(h, cudaMemsetAsync(h->table_spare, 0, h->table_cap * sizeof(Slot), st));
    CK(h, cudaMemsetAsync(h->state_new, 0, h->table_cap, st));
    CK(h, cudaMemsetAsync(h->ctr, 0, sizeof(Counters), st));
    const int grid = (int)std::min<u64>((h->table_cap + 255) / 256, (u64)h->num_sms * 16);
    rebuild_table_kernel<<<grid, 256, 0, st>>>(h->table, h->state_old, h->table_cap, h->table_spare, h->state_new,
                                               (u32)(h->table_cap - 1), h->ctr);
    CK(h, cudaGetLastError());
    Counters c;
    CK(h, cudaMemcpyAsync(&c, h->ctr, sizeof(Counters), cudaMemcpyDeviceToHost, st));
    CK(h, cudaEventRecord(e1, st));
    CK(h, cudaStreamSynchronize(st));
    cudaEventElapsedTime(&h->ms_last_rebuild, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (c.flags & kFlagTableFull) FAIL(h, PTE_ERR_TABLE_FULL, "more pods alive at once than the engine was created for (max_pods)");
    std::swap(h->table, h->table_spare);
    std::swap(h->state_old, h->state_new);
    h->pods_total = c.n_pods;
    h->n_rebuilds++;
    return PTE_OK;
}

// rv, namespace and phase of the emitted records are gathered from the log the first time somebody asks for them
int ensure_full_columns(pte_engine* h) {
    if (h->full_cols_valid) return PTE_OK;
    const u64 n = h->last.n_emitted;
    if (n) {
        int grid = (int)std::min<u64>((n + 255) / 256, (u64)h->num_sms * 16);
        emitted_columns_kernel<<<grid, 256, 0, h->stream>>>(h->out_idx, n, h->rv, h->ns, h->phase, h->out_rv, h->out_ns, h->out_phase);
        CK(h, cudaGetLastError());
    }
    h->full_cols_valid = true;
    return PTE_OK;
}

// Bring a log whose rv decreases somewhere into replay order (ascending rv, ties by arrival) in shadow
// columns, run the pipeline on those, map the emitted indices back to arrival indices.
int run_reordered(pte_engine* h, int flags, pte_stats* out) {
    const u64 n = h->n_events, c = h->cap_ev;
    cudaStream_t st = h->stream;
    if (!h->d_minmax) CK(h, cudaMalloc((void**)&h->d_minmax, 2 * sizeof(u64)));
    const u64 init[2] = {~0ull, 0ull};
    CK(h, cudaMemcpyAsync(h->d_minmax, init, sizeof(init), cudaMemcpyHostToDevice, st));
    int grid = (int)std::min<u64>((n + 255) / 256, (u64)h->num_sms * 16);
    rv_minmax_kernel<<<grid, 256, 0, st>>>(h->rv, n, h->d_minmax);
    CK(h, cudaGetLastError());
    u64 mm[2];
    CK(h, cudaMemcpyAsync(mm, h->d_minmax, sizeof(mm), cudaMemcpyDeviceToHost, st));
    CK(h, cudaStreamSynchronize(st));
    const u32 span_bits = std::max<u32>(1, bitlen_host(mm[1] - mm[0]));
    if (span_bits + h->L.idx_bits > 64)
        FAIL(h, PTE_ERR_UNORDERED, "resourceVersion decreases along arrival order and its span is too wide to pre-sort "
                                   "(bits(rv_max - rv_min) + bits(max_events) > 64): sort the log by rv before ingest");
    if (!h->perm) {
        if (dev_alloc(h, &h->shadow.uid_hi, c) || dev_alloc(h, &h->shadow.uid_lo, c) || dev_alloc(h, &h->shadow.rv, c) ||
            dev_alloc(h, &h->shadow.ns, c) || dev_alloc(h, &h->shadow.phase, c) || dev_alloc(h, &h->shadow.type, c) ||
            dev_alloc(h, &h->perm, c))
            return PTE_ERR_CUDA;
        CK(h, cudaMemsetAsync(h->shadow.rv, 0, c * sizeof(u64), st));
    }
    rv_pack_kernel<<<grid, 256, 0, st>>>(h->rv, n, mm[0], h->L.idx_bits, h->recA);
    CK(h, cudaGetLastError());
    u64* sorted = nullptr;
    h->pass_ev_used = 0;
    int rc = sort_field(h, h->recA, h->recB, h->recC, nullptr, n, (int)h->L.idx_bits, (int)(h->L.idx_bits + span_bits), &sorted);
    if (rc) return rc;
    LogColumns src{h->uid_hi, h->uid_lo, h->rv, h->ns, h->phase, h->type};
    permute_log_kernel<<<grid, 256, 0, st>>>(sorted, n, h->L.idx_mask(), src, h->shadow, h->perm);
    CK(h, cudaGetLastError());
    // run on the shadow columns
    std::swap(h->uid_hi, h->shadow.uid_hi); std::swap(h->uid_lo, h->shadow.uid_lo); std::swap(h->rv, h->shadow.rv);
    std::swap(h->ns, h->shadow.ns); std::swap(h->phase, h->shadow.phase); std::swap(h->type, h->shadow.type);
    h->in_reorder = true;
    rc = pte_run(h, flags & ~PTE_RUN_BITMAP_ONLY, out);      // the bitmap of the sorted copy is of no use to the caller: assemble columns
    h->in_reorder = false;
    std::swap(h->uid_hi, h->shadow.uid_hi); std::swap(h->uid_lo, h->shadow.uid_lo); std::swap(h->rv, h->shadow.rv);
    std::swap(h->ns, h->shadow.ns); std::swap(h->phase, h->shadow.phase); std::swap(h->type, h->shadow.type);
    if (rc) return rc;
    if (h->last.n_emitted) {
        grid = (int)std::min<u64>((h->last.n_emitted + 255) / 256, (u64)h->num_sms * 16);
        map_idx_kernel<<<grid, 256, 0, st>>>(h->out_idx, h->ctr, h->perm);
        CK(h, cudaGetLastError());
        CK(h, cudaStreamSynchronize(st));
    }
    h->last_ordered = 0;
    if (out) out->ordered = 0;
    return PTE_OK;
}

}  // namespace

extern "C" {

int pte_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

const char* pte_last_error(pte_handle h) { return h ? h->err.c_str() : g_err.c_str(); }

void pte_destroy(pte_handle h) {
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(h->comm);
    void* bufs[] = {h->uid_hi, h->uid_lo, h->rv, h->ns, h->phase, h->type, h->table, h->table_spare, h->rec_pool, h->part_hist, h->part_start, h->part_nsegs, h->segs, h->seg_export, h->bitmap, h->keepmap, h->state_old, h->state_new, h->lookback,
                    h->hist, h->ctr, h->ns_bitmap, h->out_idx, h->out_ns, h->out_rv, h->out_key, h->out_phase,
                    h->out_type, h->d_counts_send, h->d_counts_recv, h->shadow.uid_hi, h->shadow.uid_lo, h->shadow.rv,
                    h->shadow.ns, h->shadow.phase, h->shadow.type, h->perm, h->d_minmax, h->pay_buf, h->pay_blob,
                    h->pay_frag_off, h->pay_rec_off, h->pay_tile_off, h->json_text, h->json_nl, h->json_chunk_cnt, h->json_ns_keys, h->json_ns_off,
                    h->json_ph_keys, h->json_ph_off, h->json_err, h->json_ns_remap, h->json_ph_remap, h->json_uid_keys, h->json_uid_off, h->json_uid_remap,
                    h->tiny_in_dev, h->tiny_out_dev};
    for (void* b : bufs) if (b) cudaFree(b);
    if (h->tiny_in_host) cudaFreeHost(h->tiny_in_host);
    if (h->tiny_out_host) cudaFreeHost(h->tiny_out_host);
    for (auto e : h->ev) cudaEventDestroy(e);
    for (auto& pe : h->pass_ev) { cudaEventDestroy(pe.first); cudaEventDestroy(pe.second); }
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

int pte_create(int device, uint64_t max_events, uint64_t max_pods, pte_handle* out) {
    if (!out) FAIL((pte_handle) nullptr, PTE_ERR_INVALID, "out is NULL");
    *out = nullptr;
    int ndev = pte_device_count();
    if (ndev <= 0) FAIL((pte_handle) nullptr, PTE_ERR_NO_DEVICE, "no CUDA device: the engine has no CPU path");
    if (device < 0 || device >= ndev) FAIL((pte_handle) nullptr, PTE_ERR_INVALID, "device index out of range");
    if (max_events == 0) max_events = 1;
    if (max_events > kMaxEvents)
        FAIL((pte_handle) nullptr, PTE_ERR_CAPACITY, "max_events above 2^30-1 per engine: shard the log by uid");
    if (max_pods == 0) max_pods = max_events;      // the dictionary is sized by max_pods alone: a stream of short batches may know many pods

    pte_engine* h = new pte_engine();
    h->device = device; h->max_events = max_events; h->max_pods = max_pods;
    u64 want = (u64)((double)max_pods / 0.6) + 1;
    u64 cap = 256;
    while (cap < want) cap <<= 1;
    h->table_cap = cap;
    h->L.idx_bits = std::max<u32>(1, bitlen_host(max_events - 1));
    h->L.key_bits = bitlen_host(cap - 1);
    if (8 + h->L.idx_bits + h->L.key_bits > 64) {
        g_err = "record layout does not fit 64 bits (idx_bits + key_bits > 56): lower max_events or max_pods";
        delete h;
        return PTE_ERR_CAPACITY;
    }
    if (const char* ni = getenv("PTE_NO_INVERT")) h->no_invert = atoi(ni) != 0;
    if (const char* nd = getenv("PTE_NO_DEFER")) h->no_defer = atoi(nd) != 0;
    if (const char* ty = getenv("PTE_TINY")) h->tiny_on = atoi(ty) != 0;
    if (const char* sg = getenv("PTE_STAGE_TIMING")) h->force_stage_timing = atoi(sg) != 0;
    if (const char* rb = getenv("PTE_DIFF_ROUND_MB")) h->diff_round_bytes = std::max(1, atoi(rb)) * (1ull << 20);
    if (const char* sc = getenv("PTE_SCATTER_CFG")) h->scatter_cfg = std::min(std::max(atoi(sc), 0), 3);
    if (const char* pm = getenv("PTE_PART_MIN")) h->part_min_events = (u64)atof(pm);
    if (const char* sq = getenv("PTE_SEG_Q8")) h->seg_q8 = (u32)std::min(std::max(atoi(sq), 32), 4096);
    if (const char* pp = getenv("PTE_PATH")) h->path_mode = !strcmp(pp, "sort") ? 1 : (!strcmp(pp, "partition") ? 2 : 0);
    const char* cfg = getenv("PTE_PASS_CFG");
    if (cfg) h->pass_cfg = std::min(std::max(atoi(cfg), 0), kNumPassCfgs - 1);

#define CKC(call)                                                                      \
    do {                                                                               \
        int rc__ = (call);                                                             \
        if (rc__ != PTE_OK) { g_err = h->err; pte_destroy(h); return rc__; }           \
    } while (0)
#define CKCU(call)                                                                     \
    do {                                                                               \
        cudaError_t e__ = (call);                                                      \
        if (e__ != cudaSuccess) {                                                      \
            g_err = std::string(#call) + ": " + cudaGetErrorString(e__);               \
            pte_destroy(h);                                                            \
            return PTE_ERR_CUDA;                                                       \
        }                                                                              \
    } while (0)

    CKCU(cudaSetDevice(device));
    if (const char* g = getenv("PTE_L2_FETCH")) cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)atoi(g));
    CKCU(cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, device));
    CKCU(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    CKC(prepare_pass_cfg(h));
    // the partition-first path views the dictionary as kParts slices of part_slice slots (at least 256 each)
    h->part_slice = (u32)std::max<u64>(256, h->table_cap / kParts);
    h->table_alloc = std::max<u64>(h->table_cap, (u64)kParts * h->part_slice);
    CKC(dev_alloc(h, &h->table, h->table_alloc));
    CKC(dev_alloc(h, &h->state_old, h->table_cap)); CKC(dev_alloc(h, &h->state_new, h->table_cap));
    CKC(dev_alloc(h, &h->part_hist, (u64)kParts)); CKC(dev_alloc(h, &h->part_start, (u64)kParts + 1));
    CKC(dev_alloc(h, &h->part_nsegs, 2)); CKC(dev_alloc(h, &h->segs, (u64)kMaxSegments + kSampleSegments));
    CKC(prepare_partition(h));
    CKC(dev_alloc(h, &h->hist, (u64)kMaxPasses * kRadix));
    CKC(dev_alloc(h, &h->ctr, 1));
    CKCU(cudaFuncSetAttribute(tiny_replay_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTinySmem));
    {
        const size_t in_bytes = (size_t)kTinyMax * (8 + 8 + 8 + 4 + 1 + 1), out_bytes = sizeof(Counters) + (size_t)kTinyMax * 5;
        CKC(dev_alloc(h, &h->tiny_in_dev, in_bytes)); CKC(dev_alloc(h, &h->tiny_out_dev, out_bytes));
        CKCU(cudaHostAlloc((void**)&h->tiny_in_host, in_bytes, cudaHostAllocDefault));
        CKCU(cudaHostAlloc((void**)&h->tiny_out_host, out_bytes, cudaHostAllocDefault));
    }
    CKC(alloc_event_buffers(h));
    CKC(dev_alloc(h, &h->d_counts_send, 4));
    CKCU(cudaMemsetAsync(h->ctr, 0, sizeof(Counters), h->stream));
    for (int i = 0; i < 8; i++) { cudaEvent_t e; CKCU(cudaEventCreate(&e)); h->ev.push_back(e); }
    CKCU(cudaStreamSynchronize(h->stream));
#undef CKC
#undef CKCU
    *out = h;
    return PTE_OK;
}

int pte_set_filters(pte_handle h, const uint32_t* ns_allow, int n_allow, int ns_filter_on, int production_critical_only) {
    if (!h) FAIL(h, PTE_ERR_INVALID, "handle is NULL");
    if (n_allow < 0 || (n_allow > 0 && !ns_allow)) FAIL(h, PTE_ERR_INVALID, "bad namespace allow-list");
    CK(h, cudaSetDevice(h->device));
    u32 max_id = 0;
    for (int i = 0; i < n_allow; i++) max_id = std::max(max_id, ns_allow[i]);
    u32 words = n_allow ? max_id / 32 + 1 : 1;
    std::vector<u32> bm(words, 0u);
    for (int i = 0; i < n_allow; i++) bm[ns_allow[i] >> 5] |= 1u << (ns_allow[i] & 31);
    // the bitmap buffer only ever grows; the copy is ordered on the engine stream like every kernel that reads it
    if (words > h->ns_bitmap_cap) {
        if (h->ns_bitmap) { CK(h, cudaStreamSynchronize(h->stream)); CK(h, cudaFree(h->ns_bitmap)); h->ns_bitmap = nullptr; h->ns_bitmap_cap = 0; }
        const u32 cap = std::max<u32>(64, words * 2);
        CK(h, cudaMalloc((void**)&h->ns_bitmap, cap * sizeof(u32)));
        h->ns_bitmap_cap = cap;
    }
    h->ns_bitmap_host.assign(bm.begin(), bm.end());       // stays alive until the async copy has been consumed
    CK(h, cudaMemcpyAsync(h->ns_bitmap, h->ns_bitmap_host.data(), words * sizeof(u32), cudaMemcpyHostToDevice, h->stream));
    h->ns_words = words;
    h->ns_filter_on = (n_allow > 0 || ns_filter_on) ? 1 : 0;
    h->critical_only = production_critical_only ? 1 : 0;
    return PTE_OK;
}

static int ingest_impl(pte_handle h, const uint64_t* uid_hi, const uint64_t* uid_lo, const uint64_t* rv, const uint32_t* ns_id,
                       const uint8_t* phase, const uint8_t* type, uint64_t n, bool sync);

int pte_ingest(pte_handle h, const uint64_t* uid_hi, const uint64_t* uid_lo, const uint64_t* rv, const uint32_t* ns_id,
               const uint8_t* phase, const uint8_t* type, uint64_t n) {
    return ingest_impl(h, uid_hi, uid_lo, rv, ns_id, phase, type, n, true);
}

// One batch of a live stream in ONE call: copy in, replay as the next PTE_RUN_CARRY batch, arrival index + event type of
// the notifications out.  Same as pte_ingest + pte_run + pte_fetch(idx, type) with two host synchronisations less.
static int tiny_unstage(pte_handle h);
int pte_stream_batch(pte_handle h, const uint64_t* uid_hi, const uint64_t* uid_lo, const uint64_t* rv, const uint32_t* ns_id,
                     const uint8_t* phase, const uint8_t* type, uint64_t n, int flags, uint32_t* idx_out, uint8_t* type_out,
                     uint64_t cap, pte_stats* out) {
    int rc;
    if (h && h->tiny_on && n && n <= kTinyMax && n <= h->max_events && (flags & PTE_RUN_DEDUPE) && uid_hi && uid_lo && rv && phase && type &&
        (ns_id || !h->ns_filter_on)) {
        // a small batch: packed in pinned memory, ONE copy to the device; the replay kernel spreads it over the columns
        CK(h, cudaSetDevice(h->device));
        u8* q = h->tiny_in_host;
        memcpy(q, uid_hi, n * 8); q += n * 8;
        memcpy(q, uid_lo, n * 8); q += n * 8;
        memcpy(q, rv, n * 8); q += n * 8;
        if (ns_id) memcpy(q, ns_id, n * 4); else memset(q, 0, n * 4);
        q += n * 4;
        memcpy(q, phase, n); q += n;
        memcpy(q, type, n); q += n;
        CK(h, cudaMemcpyAsync(h->tiny_in_dev, h->tiny_in_host, (size_t)(q - h->tiny_in_host), cudaMemcpyHostToDevice, h->stream));
        h->tiny_staged = true;
        h->ns_loaded = true;
        h->n_events = n;
        h->has_run = false;
        h->batch_first_rv = rv[0];
        h->batch_last_rv = rv[n - 1];
    } else {
        rc = ingest_impl(h, uid_hi, uid_lo, rv, ns_id, phase, type, n, false);
        if (rc) return rc;
    }
    pte_stats st;
    rc = pte_run(h, flags & ~PTE_RUN_BITMAP_ONLY, &st);      // synchronises the stream: the host columns are free again
    if (rc) { tiny_unstage(h); return rc; }                  // (a batch refused before its replay still ends up in the columns)
    if (out) *out = st;
    if (st.n_emitted > cap) FAIL(h, PTE_ERR_CAPACITY, "buffers smaller than the emitted record count");
    if (st.n_emitted && h->tiny_result) {
        const u8* r = h->tiny_out_host + sizeof(Counters);
        if (idx_out) memcpy(idx_out, r, st.n_emitted * 4);
        if (type_out) memcpy(type_out, r + st.n_emitted * 4, st.n_emitted);
    } else if (st.n_emitted && (idx_out || type_out)) {
        if (idx_out) CK(h, cudaMemcpyAsync(idx_out, h->out_idx, st.n_emitted * 4, cudaMemcpyDeviceToHost, h->stream));
        if (type_out) CK(h, cudaMemcpyAsync(type_out, h->out_type, st.n_emitted, cudaMemcpyDeviceToHost, h->stream));
        CK(h, cudaStreamSynchronize(h->stream));
    }
    return PTE_OK;
}

static int ingest_impl(pte_handle h, const uint64_t* uid_hi, const uint64_t* uid_lo, const uint64_t* rv, const uint32_t* ns_id,
                       const uint8_t* phase, const uint8_t* type, uint64_t n, bool sync) {
    if (!h) FAIL(h, PTE_ERR_INVALID, "handle is NULL");
    if (n > h->max_events) FAIL(h, PTE_ERR_CAPACITY, "log longer than the engine's max_events");
    if (n && (!uid_hi || !uid_lo || !rv || !phase || !type)) FAIL(h, PTE_ERR_INVALID, "NULL column");
    if (n && !ns_id && h->ns_filter_on) FAIL(h, PTE_ERR_INVALID, "ns_id may only be NULL while no namespace allow-list is set");
    CK(h, cudaSetDevice(h->device));
    if (n) {
        CK(h, cudaMemcpyAsync(h->uid_hi, uid_hi, n * 8, cudaMemcpyHostToDevice, h->stream));
        CK(h, cudaMemcpyAsync(h->uid_lo, uid_lo, n * 8, cudaMemcpyHostToDevice, h->stream));
        CK(h, cudaMemcpyAsync(h->rv, rv, n * 8, cudaMemcpyHostToDevice, h->stream));
        if (ns_id) CK(h, cudaMemcpyAsync(h->ns, ns_id, n * 4, cudaMemcpyHostToDevice, h->stream));
        else if (h->ns_loaded) CK(h, cudaMemsetAsync(h->ns, 0, n * 4, h->stream));      // column absent: every event reads as namespace 0
        h->ns_loaded = ns_id != nullptr;
        CK(h, cudaMemcpyAsync(h->phase, phase, n, cudaMemcpyHostToDevice, h->stream));
        CK(h, cudaMemcpyAsync(h->type, type, n, cudaMemcpyHostToDevice, h->stream));
        if (sync) CK(h, cudaStreamSynchronize(h->stream));
    }
    h->n_events = n;
    h->has_run = false;
    h->batch_first_rv = n ? rv[0] : 0;
    h->batch_last_rv = n ? rv[n - 1] : 0;
    return PTE_OK;
}

int pte_read_log(pte_handle h, uint64_t* uid_hi, uint64_t* uid_lo, uint64_t* rv, uint32_t* ns_id, uint8_t* phase,
                 uint8_t* type, uint64_t cap, uint64_t* n_out) {
    if (!h) FAIL(h, PTE_ERR_INVALID, "handle is NULL");
    const u64 n = h->n_events;
    if (n_out) *n_out = n;
    if (cap < n) FAIL(h, PTE_ERR_CAPACITY, "buffers smaller than the log");
    CK(h, cudaSetDevice(h->device));
    if (n) {
        if (uid_hi) CK(h, cudaMemcpyAsync(uid_hi, h->uid_hi, n * 8, cudaMemcpyDeviceToHost, h->stream));
        if (uid_lo) CK(h, cudaMemcpyAsync(uid_lo, h->uid_lo, n * 8, cudaMemcpyDeviceToHost, h->stream));
        if (rv) CK(h, cudaMemcpyAsync(rv, h->rv, n * 8, cudaMemcpyDeviceToHost, h->stream));
        if (ns_id) CK(h, cudaMemcpyAsync(ns_id, h->ns, n * 4, cudaMemcpyDeviceToHost, h->stream));
        if (phase) CK(h, cudaMemcpyAsync(phase, h->phase, n, cudaMemcpyDeviceToHost, h->stream));
        if (type) CK(h, cudaMemcpyAsync(type, h->type, n, cudaMemcpyDeviceToHost, h->stream));
        CK(h, cudaStreamSynchronize(h->stream));
    }
    return PTE_OK;
}

static int check_gen(const pte_gen_params* p, std::string& why) {
    if (!p) { why = "params are NULL"; return PTE_ERR_INVALID; }
    if (p->n_shards == 0 || p->shard >= p->n_shards) { why = "shard must be in [0, n_shards)"; return PTE_ERR_INVALID; }
    if (p->n_events && p->n_pods == 0) { why = "n_pods must be > 0"; return PTE_ERR_INVALID; }
    if (p->dist > 1) { why = "dist must be 0 (uniform) or 1 (octave-Zipf)"; return PTE_ERR_INVALID; }
    if (p->n_pods >= (1ull << 32)) { why = "n_pods must be below 2^32"; return PTE_ERR_INVALID; }
    return PTE_OK;
}

static GenParams to_gen(const pte_gen_params* p) {
    GenParams g;
    g.seed = p->seed; g.n_events = p->n_events; g.n_pods = p->n_pods; g.dist = p->dist;
    g.p_del_q24 = p->p_del_q24; g.p_nostatus_q24 = p->p_nostatus_q24; g.p_null_q24 = p->p_null_q24;
    g.shard = p->shard; g.n_shards = p->n_shards;
    return g;
}

int pte_generate_host(const pte_gen_params* p, uint64_t* uid_hi, uint64_t* uid_lo, uint64_t* rv, uint32_t* ns_id,
                      uint8_t* phase, uint8_t* type) {
    std::string why;
    int rc = check_gen(p, why);
    if (rc) FAIL((pte_handle) nullptr, rc, why);
    GenParams g = to_gen(p);
    // per-pod state: 0 no event yet, 1 previous event had no delete draw, 2 it had one
    std::vector<u8> st(g.n_pods ? g.n_pods : 1, 0);
    for (u64 j = 0; j < g.n_events; j++) {
        GenEvent e = gen_event(g, j);
        u8 s = st[e.q];
        if (uid_hi) uid_hi[j] = e.uid_hi;
        if (uid_lo) uid_lo[j] = e.uid_lo;
        if (rv) rv[j] = e.rv;
        if (ns_id) ns_id[j] = e.ns;
        if (phase) phase[j] = (u8)e.phase;
        if (type) type[j] = (u8)gen_type(s != 0, s == 2, e.draw);
        st[e.q] = e.draw ? 2 : 1;
    }
    return PTE_OK;
}

int pte_generate(pte_handle h, const pte_gen_params* p) {
    if (!h) FAIL(h, PTE_ERR_INVALID, "handle is NULL");
    std::string why;
    int rc = check_gen(p, why);
    if (rc) FAIL(h, rc, why);
    if (p->n_events > h->max_events) FAIL(h, PTE_ERR_CAPACITY, "n_events above the engine's max_events");
    CK(h, cudaSetDevice(h->device));
    GenArgs a;
    a.g = to_gen(p);
    a.uid_hi = h->uid_hi; a.uid_lo = h->uid_lo; a.rv = h->rv; a.ns = h->ns; a.phase = h->phase; a.type = h->type;
    a.rec_out = h->recA;
    a.L.idx_bits = h->L.idx_bits;
    a.L.key_bits = std::max<u32>(1, bitlen_host(p->n_pods ? p->n_pods - 1 : 0));
    if (8 + a.L.idx_bits + a.L.key_bits > 64) FAIL(h, PTE_ERR_CAPACITY, "n_pods too large for the generator's sort record");
    const u64 n = p->n_events;
    h->n_events = n;
    h->has_run = false;
    h->ns_loaded = true;
    if (n == 0) return PTE_OK;
    int grid = (int)std::min<u64>((n + 255) / 256, (u64)h->num_sms * 16);
    gen_fill_kernel<<<grid, 256, 0, h->stream>>>(a);
    CK(h, cudaGetLastError());
    u64* sorted = nullptr;
    h->pass_ev_used = 0;
    rc = sort_field(h, h->recA, h->recB, h->recC, nullptr, n, a.L.key_shift(), a.L.key_shift() + a.L.key_bits, &sorted);
    if (rc) return rc;
    gen_type_kernel<<<grid, 256, 0, h->stream>>>(sorted, n, a.L, h->type);
    CK(h, cudaGetLastError());
    CK(h, cudaStreamSynchronize(h->stream));
    return PTE_OK;
}

// A batch that pte_stream_batch left in the packed staging buffer -> the column arrays (what the replay kernel of the
// small-batch path does on its way; needed when the batch takes the bulk kernels or is not replayed at all).
static int tiny_unstage(pte_handle h) {
    if (!h->tiny_staged) return PTE_OK;
    const u64 n = h->n_events;
    cudaStream_t st = h->stream;
    const u8* q = h->tiny_in_dev;
    CK(h, cudaMemcpyAsync(h->uid_hi, q, n * 8, cudaMemcpyDeviceToDevice, st)); q += n * 8;
    CK(h, cudaMemcpyAsync(h->uid_lo, q, n * 8, cudaMemcpyDeviceToDevice, st)); q += n * 8;
    CK(h, cudaMemcpyAsync(h->rv, q, n * 8, cudaMemcpyDeviceToDevice, st)); q += n * 8;
    CK(h, cudaMemcpyAsync(h->ns, q, n * 4, cudaMemcpyDeviceToDevice, st)); q += n * 4;
    CK(h, cudaMemcpyAsync(h->phase, q, n, cudaMemcpyDeviceToDevice, st)); q += n;
    CK(h, cudaMemcpyAsync(h->type, q, n, cudaMemcpyDeviceToDevice, st));
    h->tiny_staged = false;
    return PTE_OK;
}

int pte_run(pte_handle h, int flags, pte_stats* out) {
    if (!h) FAIL(h, PTE_ERR_INVALID, "handle is NULL");
    CK(h, cudaSetDevice(h->device));
    const int dedupe = flags & PTE_RUN_DEDUPE;
    const bool carry = dedupe && (flags & PTE_RUN_CARRY);
    const u64 n = h->n_events;
    if (carry && h->carry_valid && n && h->batch_first_rv < h->stream_last_rv)
        FAIL(h, PTE_ERR_UNORDERED, "this batch starts at a resourceVersion below the end of the previous batch of the stream");
    h->n_launches = 0;
    h->n_diff_rounds = 0;
    h->pass_ev_used = 0;
    h->has_run = false;
    cudaStream_t st = h->stream;
    CK(h, cudaEventRecord(h->ev[0], st));
    if (carry && h->carry_valid && h->pods_total * 100 > h->table_cap * 65) {
        // the stream's dictionary is past 65 % load: drop the dead pods first (max_pods bounds LIVE pods)
        int rb = rebuild_stream_table(h);
        if (rb) { h->carry_valid = false; return rb; }
    }
    const bool fresh = !(carry && h->carry_valid);       // start from an empty dictionary
    // large one-shot T-trace replays take the partition-first path (pte_partition.cuh); streams and short logs the sort path
    const bool use_part = dedupe && !carry && n && !h->force_sort && h->path_mode != 1 &&
                          (h->path_mode == 2 || n >= h->part_min_events);
    // batches of a few thousand events: one launch does the whole replay (pte_tiny.cuh)
    const bool use_tiny = h->tiny_on && dedupe && n && n <= kTinyMax && !use_part;
    // every event record costs a short run a microsecond or two on the device and on the host: stages are timed from 2^20 events on
    const bool stages = !use_tiny && (n >= (1ull << 20) || h->force_stage_timing || use_part);
    h->stage_timing = stages;
    if (!use_tiny) { int rcu = tiny_unstage(h); if (rcu) return rcu; }     // the bulk kernels read the column arrays
    h->tiny_result = false;
    if (fresh) {
        CK(h, cudaMemsetAsync(h->table, 0, (use_part ? (u64)kParts * h->part_slice : h->table_cap) * sizeof(Slot), st));
        if (carry) CK(h, cudaMemsetAsync(h->state_old, 0, h->table_cap, st));
        h->pods_total = 0;
    }
    if (carry && !use_tiny) CK(h, cudaMemcpyAsync(h->state_new, h->state_old, h->table_cap, cudaMemcpyDeviceToDevice, st));
    if (!use_tiny) CK(h, cudaMemsetAsync(h->ctr, 0, sizeof(Counters), st));
    const bool filtered = h->critical_only || h->ns_filter_on;
    InternArgs a;
    a.uid_hi = h->uid_hi; a.uid_lo = h->uid_lo; a.rv = h->rv; a.ns = h->ns; a.phase = h->phase; a.type = h->type;
    a.n = n; a.table = h->table; a.table_mask = (u32)(h->table_cap - 1); a.L = h->L; a.rec_out = h->recA;
    a.ns_bitmap = h->ns_bitmap; a.ns_words = h->ns_words; a.ns_filter_on = h->ns_filter_on;
    a.critical_only = h->critical_only; a.ctr = h->ctr; a.keep = h->keepmap; a.block_off = h->lookback;
    a.hist = h->hist; a.key_passes = (int)(h->L.key_bits + 7) / 8;
    const u64* n_surv_ptr = &h->ctr->n_survived;
    if (use_tiny) {
        TinyArgs t;
        const u8* q = h->tiny_in_dev;
        if (h->tiny_staged) {
            t.src_hi = (const u64*)q; q += n * 8; t.src_lo = (const u64*)q; q += n * 8; t.src_rv = (const u64*)q; q += n * 8;
            t.src_ns = (const u32*)q; q += n * 4; t.src_phase = q; q += n; t.src_type = q;
        } else {
            t.src_hi = h->uid_hi; t.src_lo = h->uid_lo; t.src_rv = h->rv; t.src_ns = h->ns; t.src_phase = h->phase; t.src_type = h->type;
        }
        t.uid_hi = h->uid_hi; t.uid_lo = h->uid_lo; t.rv = h->rv; t.ns = h->ns; t.phase = h->phase; t.type = h->type;
        t.n = (u32)n; t.table = h->table; t.table_mask = (u32)(h->table_cap - 1);
        t.ns_bitmap = h->ns_bitmap; t.ns_words = h->ns_words; t.ns_filter_on = h->ns_filter_on; t.critical_only = h->critical_only;
        t.carry = carry ? 1 : 0; t.state = h->state_old; t.prev_rv = 0;
        t.ctr = h->ctr; t.bitmap = h->bitmap; t.out_idx = h->out_idx; t.out_type = h->out_type; t.packed = h->tiny_out_dev;
        tiny_replay_kernel<<<1, kTinyThreads, kTinySmem, st>>>(t);
        CK(h, cudaGetLastError());
        h->tiny_staged = false;          // the kernel has spread the batch over the column arrays
        h->n_launches = 1;
    } else if (use_part) {
        int rc = run_partitioned(h, a, filtered);
        if (rc) return rc;
    } else if (n && dedupe) {
        CK(h, cudaMemsetAsync(h->hist, 0, sizeof(u32) * kMaxPasses * kRadix, st));
        const u64 tiles = (n + kInternTile - 1) / kInternTile;
        if (filtered) {
            // predicates -> keep bitmap -> one output offset per 8192-event block -> compacting intern
            const u64 blocks = (n + kAsmTile - 1) / kAsmTile;
            int fgrid = (int)std::min<u64>((n / 4 + 255) / 256 + 1, (u64)h->num_sms * 16);
            filter_flag_kernel<<<fgrid, 256, 0, st>>>(a, h->keepmap);
            CK(h, cudaGetLastError());
            bitmap_count_kernel<<<(int)std::min<u64>(blocks, (u64)h->num_sms * 8), 256, 0, st>>>(h->keepmap, n, h->lookback);
            CK(h, cudaGetLastError());
            tile_scan_kernel<<<1, 1024, 0, st>>>(h->lookback, n, h->ctr);
            CK(h, cudaGetLastError());
            int grid = (int)std::min<u64>(blocks, (u64)h->num_sms * 8);
            intern_pack_kernel<true><<<grid, kInternThreads, 0, st>>>(a);
            h->n_launches += 3;
        } else {
            int grid = (int)std::min<u64>(tiles, (u64)h->num_sms * PTE_INTERN_BLOCKS);
            intern_pack_kernel<false><<<grid, kInternThreads, 0, st>>>(a);
        }
        CK(h, cudaGetLastError());
        h->n_launches++;
    }
    if (!use_part && stages) CK(h, cudaEventRecord(h->ev[1], st));
    AssembleArgs as;
    as.n = n; as.bitmap = h->bitmap;
    as.rv = h->rv; as.ns = h->ns; as.phase = h->phase; as.type = h->type;
    as.ctr = h->ctr; as.lookback = h->lookback;
    as.out_idx = h->out_idx; as.out_rv = h->out_rv; as.out_ns = h->out_ns;
    as.out_phase = h->out_phase; as.out_type = h->out_type;
    if (use_tiny) {
        // one kernel: only the total is timed (every event record and every elapsed-time query costs the live loop a microsecond or two)
    } else if (use_part) {
        // run_partitioned has filled the emit bitmap and recorded ev[1], ev[2]
    } else if (n && dedupe) {
        u64* sorted = nullptr;
        int rc = sort_field(h, h->recA, h->recB, h->recC, n_surv_ptr, n, h->L.key_shift(),
                            h->L.key_shift() + h->L.key_bits, &sorted, /*hist_ready=*/true);
        if (rc) return rc;
        if (stages) CK(h, cudaEventRecord(h->ev[2], st));
        const u64 per_cta = (u64)kDiffThreads * kDiffPerThread;
        int grid = (int)std::min<u64>((n + per_cta - 1) / per_cta, (u64)h->num_sms * 16);
        // polarity: flip the bits of the minority (emitted or suppressed), estimated from a sample.  Stream
        // batches keep the plain polarity (their segment heads depend on the carried state, not on the record).
        const u64 sample = 4u << 20;
        diff_sample_kernel<<<(int)std::min<u64>((std::min(n, sample) + 255) / 256, (u64)h->num_sms * 8), 256, 0, st>>>(sorted, n_surv_ptr, h->L, h->ctr, sample);
        CK(h, cudaGetLastError());
        bitmap_init_kernel<<<(int)std::min<u64>(((n + 31) / 32 + 255) / 256, (u64)h->num_sms * 8), 256, 0, st>>>(
            h->bitmap, filtered ? h->keepmap : nullptr, n, h->ctr, (carry || h->no_invert) ? 0 : 1);
        CK(h, cudaGetLastError());
        h->n_launches += 2;
        // slices: each covers a part of the bitmap small enough to stay in L2 while it is being flipped
        const u64 bits_per_round = h->diff_round_bytes * 8;
        const u64 drounds = std::max<u64>(1, (n + bits_per_round - 1) / bits_per_round);
        const u64 span = round_up((n + drounds - 1) / drounds, 32);
        DiffArgs d;
        d.in = sorted; d.n_ptr = n_surv_ptr; d.bitmap = h->bitmap; d.L = h->L; d.ctr = h->ctr;
        d.state_old = carry ? h->state_old : nullptr; d.state_new = carry ? h->state_new : nullptr;
        d.defer_list = nullptr; d.defer_cnt = nullptr; d.region_cap = 0; d.defer_hi = 0;
        // two slices: one read of the records does both -- the second slice's flips wait in per-CTA lists
        // (in the record buffer the sort has left free) until the first slice is done with the L2
        const u64 per_warp = 32ull * kDiffPerThread;
        const u64 n_lists = (u64)grid * (kDiffThreads / 32);
        const u64 region_cap = ((n + (u64)grid * per_cta - 1) / ((u64)grid * per_cta)) * per_warp;   // records one warp sees
        const bool defer = drounds == 2 && !h->no_defer && n_lists * region_cap * sizeof(u32) <= h->cap_ev * sizeof(u64) &&
                           n_lists <= h->lookback_words;
        if (defer) {
            d.pos_lo = 0; d.pos_hi = (u32)std::min(n, span); d.defer_hi = (u32)n;
            d.defer_list = reinterpret_cast<u32*>(sorted == h->recB ? h->recC : h->recB);
            d.defer_cnt = h->lookback; d.region_cap = region_cap;
            if (carry) diff_flag_kernel<true, true><<<grid, kDiffThreads, 0, st>>>(d);
            else diff_flag_kernel<false, true><<<grid, kDiffThreads, 0, st>>>(d);
            CK(h, cudaGetLastError());
            diff_apply_kernel<<<grid, 256, 0, st>>>(d.defer_list, d.defer_cnt, region_cap, (u32)n_lists, h->bitmap);
            CK(h, cudaGetLastError());
            h->n_launches += 2;
            h->n_diff_rounds = 1;
        } else {
            h->n_diff_rounds = (u32)drounds;
            for (u64 r = 0; r < drounds; r++) {
                d.pos_lo = (u32)(r * span); d.pos_hi = (u32)std::min(n, (r + 1) * span);
                if (d.pos_lo >= d.pos_hi) break;
                if (carry) diff_flag_kernel<true, false><<<grid, kDiffThreads, 0, st>>>(d);
                else diff_flag_kernel<false, false><<<grid, kDiffThreads, 0, st>>>(d);
                CK(h, cudaGetLastError());
                h->n_launches++;
            }
        }
    } else {
        // R-trace: every survivor is emitted; nothing to intern or sort
        if (stages) CK(h, cudaEventRecord(h->ev[2], st));
        if (n) {
            int grid = (int)std::min<u64>((n / 4 + 255) / 256 + 1, (u64)h->num_sms * 16);
            filter_flag_kernel<<<grid, 256, 0, st>>>(a, h->bitmap);
            CK(h, cudaGetLastError());
            h->n_launches++;
        }
    }
    if (stages) {
        CK(h, cudaEventRecord(h->ev[3], st));
        CK(h, cudaEventRecord(h->ev[4], st));
    }
    if (n && !use_tiny) {
        const u64 tiles = (n + kAsmTile - 1) / kAsmTile;
        int grid = (int)std::min<u64>(tiles, (u64)h->num_sms * 8);
        bitmap_count_kernel<<<grid, 256, 0, st>>>(h->bitmap, n, h->lookback);
        CK(h, cudaGetLastError());
        tile_scan_kernel<<<1, 1024, 0, st>>>(h->lookback, n, h->ctr);
        CK(h, cudaGetLastError());
        if (!(flags & PTE_RUN_BITMAP_ONLY)) {
            grid = (int)std::min<u64>(tiles, (u64)h->num_sms * PTE_ASM_BLOCKS);
            assemble_bitmap_kernel<<<grid, 256, 0, st>>>(as);
            CK(h, cudaGetLastError());
            h->n_launches++;
        }
        h->n_launches += 2;
    }
    CK(h, cudaEventRecord(h->ev[5], st));
    if (use_tiny) {
        // counters and notifications in one copy (at most n of them; how many is only known afterwards)
        CK(h, cudaMemcpyAsync(h->tiny_out_host, h->tiny_out_dev, sizeof(Counters) + n * 5, cudaMemcpyDeviceToHost, st));
        CK(h, cudaStreamSynchronize(st));
        memcpy(&h->last, h->tiny_out_host, sizeof(Counters));
        h->tiny_result = true;
    } else {
        CK(h, cudaMemcpyAsync(&h->last, h->ctr, sizeof(Counters), cudaMemcpyDeviceToHost, st));
        CK(h, cudaStreamSynchronize(st));
    }
    if (dedupe) h->last.n_tombstones = h->last.n_del_survived;      // every DELETED survivor is a tombstone of the T-trace
    if (use_part && (h->last.flags & (kFlagTableFull | kFlagPendingFull | kFlagSegmentsFull)) && !(h->last.flags & kFlagBadEvent)) {
        // a partition's dictionary slice, the segment table or the pending list overflowed (hash imbalance, extreme
        // skew): the sort path has no such limits
        h->force_sort = true;
        int rc2 = pte_run(h, flags, out);
        h->force_sort = false;
        return rc2;
    }
    h->last_path = use_part ? 1 : 0;
    if (carry && h->carry_valid && (h->last.flags & kFlagTableFull) && !(h->last.flags & kFlagBadEvent) && !h->retry_after_rebuild) {
        // the batch brought more new uids than the dictionary had room for: make room (dead pods go) and replay the batch.
        // Nothing of the failed attempt is kept but dictionary entries without state, which the rebuild drops.
        int rb = rebuild_stream_table(h);
        if (rb) { h->carry_valid = false; return rb; }
        h->retry_after_rebuild = true;
        int rc2 = pte_run(h, flags, out);
        h->retry_after_rebuild = false;
        return rc2;
    }
    int rc = check_flags(h, h->last.flags);
    if (rc) { h->carry_valid = false; return rc; }
    if (h->last.flags & kFlagUnordered) {
        h->carry_valid = false;
        if (carry) FAIL(h, PTE_ERR_UNORDERED, "resourceVersion decreases inside a stream batch: batches of a stream must be rv-ordered");
        if (h->in_reorder) FAIL(h, PTE_ERR_UNORDERED, "internal: the rv pre-sort did not order the log");
        return run_reordered(h, flags, out);
    }
    h->last_ordered = 1;
    h->pods_total += h->last.n_pods;
    if (carry) { if (!use_tiny) std::swap(h->state_old, h->state_new); h->stream_last_rv = n ? h->batch_last_rv : h->stream_last_rv; }
    h->carry_valid = carry;
    h->has_run = true;
    h->cols_valid = !(flags & PTE_RUN_BITMAP_ONLY);
    h->key_valid = false;
    h->full_cols_valid = false;
    if (out) {
        memset(out, 0, sizeof(*out));
        out->n_in = n; out->n_survived = h->last.n_survived; out->n_emitted = h->last.n_emitted;
        out->n_tombstones = h->last.n_tombstones; out->n_pods = h->pods_total;
        out->ordered = 1; out->n_launches = h->n_launches;
        cudaEventElapsedTime(&out->ms_total, h->ev[0], h->ev[5]);
        if (stages) {
            cudaEventElapsedTime(&out->ms_intern, h->ev[0], h->ev[1]);
            cudaEventElapsedTime(&out->ms_sort_key, h->ev[1], h->ev[2]);
            cudaEventElapsedTime(&out->ms_diff, h->ev[2], h->ev[3]);
            cudaEventElapsedTime(&out->ms_sort_order, h->ev[3], h->ev[4]);
            cudaEventElapsedTime(&out->ms_assemble, h->ev[4], h->ev[5]);
        }
        float sum = 0;
        for (size_t i = 0; stages && i < h->pass_ev_used; i++) {
            float ms = 0;
            cudaEventElapsedTime(&ms, h->pass_ev[i].first, h->pass_ev[i].second);
            sum += ms;
        }
        out->n_passes = (u32)h->pass_ev_used;
        out->n_diff_rounds = h->n_diff_rounds;
        out->ms_pass_avg = h->pass_ev_used ? sum / h->pass_ev_used : 0.f;
        out->path = (u32)h->last_path;
        out->n_rebuilds = h->n_rebuilds; out->ms_rebuild = h->ms_last_rebuild;
        if (use_part) {
            cudaEventElapsedTime(&out->ms_part_hist, h->ev[0], h->ev[6]);
            cudaEventElapsedTime(&out->ms_part_scatter, h->ev[6], h->ev[1]);
            cudaEventElapsedTime(&out->ms_part_replay, h->ev[7], h->ev[3]);
            out->n_pending = h->last.n_pending;
        }
    }
    return PTE_OK;
}

int pte_fetch(pte_handle h, uint32_t* idx, uint64_t* rv, uint64_t* key, uint32_t* ns_id, uint8_t* phase, uint8_t* type,
              uint64_t cap, uint64_t* n_out) {
    if (!h) FAIL(h, PTE_ERR_INVALID, "handle is NULL");
    if (!h->has_run) FAIL(h, PTE_ERR_STATE, "pte_fetch before a successful pte_run");
    if (!h->cols_valid) FAIL(h, PTE_ERR_STATE, "pte_fetch after a PTE_RUN_BITMAP_ONLY run: only pte_fetch_bitmap is available");
    const u64 n = h->last.n_emitted;
    if (n_out) *n_out = n;
    if (cap < n) FAIL(h, PTE_ERR_CAPACITY, "buffers smaller than the emitted record count");
    CK(h, cudaSetDevice(h->device));
    if (n) {
        cudaStream_t st = h->stream;
        if (rv || ns_id || phase) { int rc = ensure_full_columns(h); if (rc) return rc; }
        if (idx) CK(h, cudaMemcpyAsync(idx, h->out_idx, n * 4, cudaMemcpyDeviceToHost, st));
        if (rv) CK(h, cudaMemcpyAsync(rv, h->out_rv, n * 8, cudaMemcpyDeviceToHost, st));
        if (key) {
            if (!h->key_valid) {
                int grid = (int)std::min<u64>((n + 255) / 256, (u64)h->num_sms * 16);
                if (h->last_path == 1)
                    key_lookup_part_kernel<<<grid, 256, 0, st>>>(h->out_idx, n, h->uid_hi, h->uid_lo, h->table, h->part_slice, h->out_key);
                else
                    key_lookup_kernel<<<grid, 256, 0, st>>>(h->out_idx, n, h->uid_hi, h->uid_lo, h->table,
                                                            (u32)(h->table_cap - 1), h->out_key);
                CK(h, cudaGetLastError());
                h->key_valid = true;
            }
            CK(h, cudaMemcpyAsync(key, h->out_key, n * 8, cudaMemcpyDeviceToHost, st));
        }
        if (ns_id) CK(h, cudaMemcpyAsync(ns_id, h->out_ns, n * 4, cudaMemcpyDeviceToHost, st));
        if (phase) CK(h, cudaMemcpyAsync(phase, h->out_phase, n, cudaMemcpyDeviceToHost, st));
        if (type) CK(h, cudaMemcpyAsync(type, h->out_type, n, cudaMemcpyDeviceToHost, st));
        CK(h, cudaStreamSynchronize(st));
    }
    return PTE_OK;
}

int pte_fetch_bitmap(pte_handle h, uint32_t* words, uint64_t cap_words, uint64_t* n_emitted) {
    if (!h) FAIL(h, PTE_ERR_INVALID, "handle is NULL");
    if (!h->has_run) FAIL(h, PTE_ERR_STATE, "pte_fetch_bitmap before a successful pte_run");
    if (!h->last_ordered) FAIL(h, PTE_ERR_STATE, "the log was pre-sorted by rv: its emit bitmap is not in arrival order, use pte_fetch");
    const u64 need = (h->n_events + 31) / 32;
    if (n_emitted) *n_emitted = h->last.n_emitted;
    if (!words) return PTE_OK;
    if (cap_words < need) FAIL(h, PTE_ERR_CAPACITY, "bitmap buffer smaller than (n_events + 31) / 32 words");
    CK(h, cudaSetDevice(h->device));
    if (need) {
        CK(h, cudaMemcpyAsync(words, h->bitmap, need * sizeof(u32), cudaMemcpyDeviceToHost, h->stream));
        CK(h, cudaStreamSynchronize(h->stream));
    }
    return PTE_OK;
}

int pte_grow(pte_handle h, uint64_t max_events) {
    if (!h) FAIL(h, PTE_ERR_INVALID, "handle is NULL");
    if (max_events <= h->max_events) return PTE_OK;
    if (max_events > kMaxEvents) FAIL(h, PTE_ERR_CAPACITY, "max_events above 2^30-1 per engine: shard the log by uid");
    const u32 idx_bits = std::max<u32>(1, bitlen_host(max_events - 1));
    if (8 + idx_bits + h->L.key_bits > 64) FAIL(h, PTE_ERR_CAPACITY, "record layout does not fit 64 bits (idx_bits + key_bits > 56)");
    CK(h, cudaSetDevice(h->device));
    CK(h, cudaStreamSynchronize(h->stream));
    void* bufs[] = {h->uid_hi, h->uid_lo, h->rv, h->ns, h->phase, h->type, h->rec_pool, h->bitmap, h->keepmap, h->lookback,
                    h->out_idx, h->out_ns, h->out_rv, h->out_key, h->out_phase, h->out_type,
                    h->shadow.uid_hi, h->shadow.uid_lo, h->shadow.rv, h->shadow.ns, h->shadow.phase, h->shadow.type, h->perm};
    for (void* b : bufs) if (b) cudaFree(b);
    h->uid_hi = h->uid_lo = h->rv = nullptr; h->ns = nullptr; h->phase = h->type = nullptr; h->rec_pool = nullptr;
    h->bitmap = h->keepmap = h->lookback = nullptr; h->out_idx = h->out_ns = nullptr; h->out_rv = h->out_key = nullptr;
    h->out_phase = h->out_type = nullptr; h->shadow = LogColumns{}; h->perm = nullptr;
    h->max_events = max_events;
    h->L.idx_bits = idx_bits;
    h->n_events = 0; h->has_run = false; h->key_valid = false; h->ns_loaded = true;
    int rc = alloc_event_buffers(h);
    if (rc) return rc;
    CK(h, cudaStreamSynchronize(h->stream));
    return PTE_OK;
}

int pte_replay_pipelined(pte_handle h, const uint64_t* uid_hi, const uint64_t* uid_lo, const uint64_t* rv, const uint32_t* ns_id,
                         const uint8_t* phase, const uint8_t* type, uint64_t n, uint64_t chunk_events, int dedupe,
                         uint32_t* bitmap_out, pte_stats* out) {
    if (!h) FAIL(h, PTE_ERR_INVALID, "handle is NULL");
    if (n > h->max_events) FAIL(h, PTE_ERR_CAPACITY, "log longer than the engine's max_events");
    if (n && (!uid_hi || !uid_lo || !phase || !bitmap_out)) FAIL(h, PTE_ERR_INVALID, "NULL column or bitmap");
    if (n && !ns_id && h->ns_filter_on) FAIL(h, PTE_ERR_INVALID, "ns_id may only be NULL while no namespace allow-list is set");
    CK(h, cudaSetDevice(h->device));
    if (!h->copy_stream) CK(h, cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    if (chunk_events == 0) chunk_events = 1ull << 26;
    chunk_events = round_up(chunk_events, kPad);
    const u64 n_chunks = n ? (n + chunk_events - 1) / chunk_events : 0;
    pte_stats total;
    memset(&total, 0, sizeof(total));
    total.ordered = 1;
    // every host -> device copy is queued now, on its own stream; chunk k's replay waits for chunk k's event only
    std::vector<cudaEvent_t> landed(n_chunks);
    cudaStream_t cs = h->copy_stream;
    CK(h, cudaStreamSynchronize(h->stream));
    if (!ns_id && h->ns_loaded && n) CK(h, cudaMemsetAsync(h->ns, 0, n * 4, cs));
    for (u64 k = 0; k < n_chunks; k++) {
        const u64 off = k * chunk_events, len = std::min<u64>(chunk_events, n - off);
        CK(h, cudaMemcpyAsync(h->uid_hi + off, uid_hi + off, len * 8, cudaMemcpyHostToDevice, cs));
        CK(h, cudaMemcpyAsync(h->uid_lo + off, uid_lo + off, len * 8, cudaMemcpyHostToDevice, cs));
        if (rv) CK(h, cudaMemcpyAsync(h->rv + off, rv + off, len * 8, cudaMemcpyHostToDevice, cs));
        if (ns_id) CK(h, cudaMemcpyAsync(h->ns + off, ns_id + off, len * 4, cudaMemcpyHostToDevice, cs));
        CK(h, cudaMemcpyAsync(h->phase + off, phase + off, len, cudaMemcpyHostToDevice, cs));
        if (type) CK(h, cudaMemcpyAsync(h->type + off, type + off, len, cudaMemcpyHostToDevice, cs));
        else {
            // one byte per event for both: type << 6 | phase, taken apart where it lands (still on the copy stream)
            unpack_meta_kernel<<<(int)std::min<u64>((len / 16 + 255) / 256 + 1, (u64)h->num_sms * 8), 256, 0, cs>>>(h->phase + off, h->type + off, len);
            CK(h, cudaGetLastError());
        }
        CK(h, cudaEventCreateWithFlags(&landed[k], cudaEventDisableTiming));
        CK(h, cudaEventRecord(landed[k], cs));
    }
    h->ns_loaded = ns_id != nullptr;
    // the chunks are the batches of one stream (PTE_RUN_CARRY): the same notifications as one replay of the whole log
    struct Saved { u64 *uid_hi, *uid_lo, *rv; u32* ns; u8 *phase, *type; u32 *bitmap, *keepmap; u64 n_events; } sv =
        {h->uid_hi, h->uid_lo, h->rv, h->ns, h->phase, h->type, h->bitmap, h->keepmap, h->n_events};
    h->carry_valid = false;
    int rc = PTE_OK;
    const int flags = (dedupe ? (PTE_RUN_DEDUPE | PTE_RUN_CARRY) : 0) | PTE_RUN_BITMAP_ONLY;
    for (u64 k = 0; k < n_chunks && rc == PTE_OK; k++) {
        const u64 off = k * chunk_events, len = std::min<u64>(chunk_events, n - off);
        cudaStreamWaitEvent(h->stream, landed[k], 0);
        h->uid_hi = sv.uid_hi + off; h->uid_lo = sv.uid_lo + off; h->rv = sv.rv + off; h->ns = sv.ns + off;
        h->phase = sv.phase + off; h->type = sv.type + off; h->bitmap = sv.bitmap + off / 32; h->keepmap = sv.keepmap + off / 32;
        h->n_events = len;
        if (rv) {
            h->batch_first_rv = rv[off]; h->batch_last_rv = rv[off + len - 1];
            if (!dedupe && k > 0 && rv[off] < rv[off - 1]) { h->err = "resourceVersion decreases between two chunks"; rc = PTE_ERR_UNORDERED; break; }
        } else {
            // the caller vouches for arrival order: the order key is the position (the reference's loop has no other)
            iota_rv_kernel<<<(int)std::min<u64>((len + 255) / 256, (u64)h->num_sms * 8), 256, 0, h->stream>>>(h->rv, len, off + 1);
            h->batch_first_rv = off + 1; h->batch_last_rv = off + len;
        }
        pte_stats st;
        h->in_reorder = true;                                   // an unordered chunk is an error here, not a reason to pre-sort
        rc = pte_run(h, flags, &st);
        h->in_reorder = false;
        if (rc == PTE_OK) {
            cudaMemcpyAsync(bitmap_out + off / 32, h->bitmap, (len + 31) / 32 * sizeof(u32), cudaMemcpyDeviceToHost, h->stream);
            total.n_in += st.n_in; total.n_survived += st.n_survived; total.n_emitted += st.n_emitted;
            total.n_tombstones += st.n_tombstones; total.n_pods = st.n_pods; total.n_launches += st.n_launches;
            total.ms_total += st.ms_total; total.ms_intern += st.ms_intern; total.ms_sort_key += st.ms_sort_key;
            total.ms_diff += st.ms_diff; total.ms_assemble += st.ms_assemble; total.n_passes += st.n_passes;
            total.n_rebuilds = st.n_rebuilds; total.ms_rebuild = st.ms_rebuild;
        }
    }
    cudaStreamSynchronize(h->stream);
    cudaStreamSynchronize(cs);
    for (auto e : landed) if (e) cudaEventDestroy(e);
    h->uid_hi = sv.uid_hi; h->uid_lo = sv.uid_lo; h->rv = sv.rv; h->ns = sv.ns; h->phase = sv.phase; h->type = sv.type;
    h->bitmap = sv.bitmap; h->keepmap = sv.keepmap;
    h->carry_valid = false;                                     // one replay, not a stream to be continued
    h->has_run = false;                                         // the per-chunk results were handed out already
    h->n_events = rc == PTE_OK ? n : 0;
    if (rc != PTE_OK) {
        if (rc == PTE_ERR_UNORDERED)
            h->err = "pte_replay_pipelined needs a log in resourceVersion order (use pte_ingest + pte_run, which pre-sorts): " + h->err;
        return rc;
    }
    if (out) *out = total;
    return PTE_OK;
}

int pte_digest(pte_handle h, uint64_t* out) {
    if (!h || !out) FAIL(h, PTE_ERR_INVALID, "NULL argument");
    if (!h->has_run) FAIL(h, PTE_ERR_STATE, "pte_digest before a successful pte_run");
    if (!h->cols_valid) FAIL(h, PTE_ERR_STATE, "pte_digest after a PTE_RUN_BITMAP_ONLY run: only pte_fetch_bitmap is available");
    CK(h, cudaSetDevice(h->device));
    CK(h, cudaMemsetAsync(&h->ctr->digest, 0, sizeof(u64), h->stream));
    const u64 n = h->last.n_emitted;
    if (n) {
        int grid = (int)std::min<u64>((n + 255) / 256, (u64)h->num_sms * 16);
        digest_kernel<<<grid, 256, 0, h->stream>>>(h->out_idx, h->phase, h->out_type, h->uid_hi, h->uid_lo, h->rv,
                                                    h->ns, h->ctr);
        CK(h, cudaGetLastError());
    }
    u64 d = 0;
    CK(h, cudaMemcpyAsync(&d, &h->ctr->digest, sizeof(u64), cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    *out = d;
    return PTE_OK;
}

int pte_serialize_payloads(pte_handle h, uint64_t first, uint64_t count, const char* blob, uint32_t blob_len,
                           const uint32_t* frag_off, uint32_t n_phase, uint32_t n_ns, char* out, uint64_t out_cap,
                           uint64_t* offsets, uint64_t* n_bytes) {
    if (!h || !blob || !frag_off || !n_bytes) FAIL(h, PTE_ERR_INVALID, "NULL argument");
    if (!h->has_run) FAIL(h, PTE_ERR_STATE, "pte_serialize_payloads before a successful pte_run");
    if (!h->cols_valid) FAIL(h, PTE_ERR_STATE, "pte_serialize_payloads after a PTE_RUN_BITMAP_ONLY run: only pte_fetch_bitmap is available");
    if (first > h->last.n_emitted || count > h->last.n_emitted - first) FAIL(h, PTE_ERR_INVALID, "record range outside the emitted records");
    if (n_phase == 0 || n_ns == 0) FAIL(h, PTE_ERR_INVALID, "empty phase or namespace table");
    *n_bytes = 0;
    if (count == 0) { if (offsets) offsets[0] = 0; return PTE_OK; }
    CK(h, cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    { int rcf = ensure_full_columns(h); if (rcf) return rcf; }
    const u64 n_frag = 9ull + n_phase + n_ns;
    if (frag_off[n_frag] != blob_len) FAIL(h, PTE_ERR_INVALID, "fragment offsets do not end at blob_len");
    int rc;
    if ((rc = grow(h, &h->pay_blob, &h->pay_blob_cap, (u64)blob_len + 16))) return rc;
    if ((rc = grow(h, &h->pay_frag_off, &h->pay_frag_cap, n_frag + 1))) return rc;
    const u64 tiles = (count + kPayTile - 1) / kPayTile;
    if ((rc = grow(h, &h->pay_tile_off, &h->pay_tile_cap, tiles + 1))) return rc;
    if ((rc = grow(h, &h->pay_rec_off, &h->pay_off_cap, count + 1))) return rc;
    if (tiles > h->lookback_words) FAIL(h, PTE_ERR_CAPACITY, "too many records in one call");
    CK(h, cudaMemcpyAsync(h->pay_blob, blob, blob_len, cudaMemcpyHostToDevice, st));
    CK(h, cudaMemcpyAsync(h->pay_frag_off, frag_off, (n_frag + 1) * sizeof(u32), cudaMemcpyHostToDevice, st));
    PayloadTable t{h->pay_blob, h->pay_frag_off, n_phase, n_ns};
    int grid = (int)std::min<u64>(tiles, (u64)h->num_sms * 8);
    payload_len_kernel<<<grid, 256, 0, st>>>(t, h->out_ns, h->out_phase, h->out_type, first, count, h->lookback);
    CK(h, cudaGetLastError());
    scan_u32_to_u64_kernel<<<1, 1024, 0, st>>>(h->lookback, tiles, h->pay_tile_off);
    CK(h, cudaGetLastError());
    u64 total = 0;
    CK(h, cudaMemcpyAsync(&total, h->pay_tile_off + tiles, sizeof(u64), cudaMemcpyDeviceToHost, st));
    CK(h, cudaStreamSynchronize(st));
    *n_bytes = total;
    if (!out || !offsets) return PTE_OK;                       // size query
    if (total > out_cap) FAIL(h, PTE_ERR_CAPACITY, "output buffer smaller than the serialised payloads (size in *n_bytes)");
    if ((rc = grow(h, &h->pay_buf, &h->pay_buf_cap, total + 64))) return rc;
    CK(h, cudaEventRecord(h->ev[3], st));
    payload_write_kernel<<<grid, 256, 0, st>>>(t, h->out_idx, h->out_ns, h->out_phase, h->out_type, h->uid_hi, h->uid_lo,
                                                first, count, h->pay_tile_off, h->pay_buf, h->pay_rec_off);
    CK(h, cudaGetLastError());
    CK(h, cudaEventRecord(h->ev[4], st));
    CK(h, cudaMemcpyAsync(out, h->pay_buf, total, cudaMemcpyDeviceToHost, st));
    CK(h, cudaMemcpyAsync(offsets, h->pay_rec_off, count * sizeof(u64), cudaMemcpyDeviceToHost, st));
    CK(h, cudaStreamSynchronize(st));
    offsets[count] = total;
    cudaEventElapsedTime(&h->ms_payload, h->ev[3], h->ev[4]);
    h->payload_bytes = total;
    return PTE_OK;
}

static constexpr u32 kJsonNsSlots = 1u << 16, kJsonPhSlots = 128, kJsonUidSlots = 1u << 20;      // uid strings that are not canonical UUIDs

struct HostJsonText { const char* base; __host__ __device__ char operator[](u32 i) const { return base[i]; } };

int pte_ingest_json(pte_handle h, const char* text, uint64_t len, uint64_t* n_events, uint64_t* bad_line) {
    if (!h || (!text && len)) FAIL(h, PTE_ERR_INVALID, "NULL argument");
    if (n_events) *n_events = 0;
    if (bad_line) *bad_line = 0;
    if (len >= (1ull << 32) - 64) FAIL(h, PTE_ERR_CAPACITY, "watch log of 4 GiB or more: ingest it in pieces");
    CK(h, cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    h->has_run = false;
    if (!h->json_stream) { for (auto& v : h->json_names) v.clear(); }
    if (len == 0) { h->n_events = 0; return PTE_OK; }
    int rc;
    if ((rc = grow(h, &h->json_text, &h->json_text_cap, len + 64))) return rc;
    CK(h, cudaMemcpyAsync(h->json_text, text, len, cudaMemcpyHostToDevice, st));
    CK(h, cudaMemsetAsync(h->json_text + len, 0, 64, st));                    // the kernels read whole 16- and 32-byte blocks
    // 1. line index
    const u64 chunks = (len + kJsonChunk - 1) / kJsonChunk;
    if ((rc = grow(h, &h->json_chunk_cnt, &h->json_chunk_cap, chunks))) return rc;
    if ((rc = grow(h, &h->pay_tile_off, &h->pay_tile_cap, chunks + 1))) return rc;
    int grid = (int)std::min<u64>(chunks, (u64)h->num_sms * 8);
    CK(h, cudaEventRecord(h->ev[0], st));
    json_count_newlines_kernel<<<grid, 256, 0, st>>>(h->json_text, len, h->json_chunk_cnt);
    CK(h, cudaGetLastError());
    scan_u32_to_u64_kernel<<<1, 1024, 0, st>>>(h->json_chunk_cnt, chunks, h->pay_tile_off);
    CK(h, cudaGetLastError());
    u64 n_nl = 0;
    CK(h, cudaMemcpyAsync(&n_nl, h->pay_tile_off + chunks, sizeof(u64), cudaMemcpyDeviceToHost, st));
    CK(h, cudaStreamSynchronize(st));
    const u64 n_lines = n_nl + (text[len - 1] != '\n' ? 1 : 0);
    if (n_lines > h->max_events) FAIL(h, PTE_ERR_CAPACITY, "more watch events than the engine's max_events");
    if ((rc = grow(h, &h->json_nl, &h->json_nl_cap, n_nl + 1))) return rc;
    json_list_newlines_kernel<<<grid, 256, 0, st>>>(h->json_text, len, h->pay_tile_off, h->json_nl);
    CK(h, cudaGetLastError());
    // 2. dictionaries + parse
    if (!h->json_ns_keys) {
        CK(h, cudaMalloc((void**)&h->json_ns_keys, kJsonNsSlots * sizeof(u64)));
        CK(h, cudaMalloc((void**)&h->json_ns_off, kJsonNsSlots * sizeof(u32)));
        CK(h, cudaMalloc((void**)&h->json_ph_keys, kJsonPhSlots * sizeof(u64)));
        CK(h, cudaMalloc((void**)&h->json_ph_off, kJsonPhSlots * sizeof(u32)));
        CK(h, cudaMalloc((void**)&h->json_err, 4 * sizeof(u32)));
        CK(h, cudaMalloc((void**)&h->json_ns_remap, kJsonNsSlots * sizeof(u32)));
        CK(h, cudaMalloc((void**)&h->json_ph_remap, kJsonPhSlots));
        CK(h, cudaMalloc((void**)&h->json_uid_keys, kJsonUidSlots * sizeof(u64)));
        CK(h, cudaMalloc((void**)&h->json_uid_off, kJsonUidSlots * sizeof(u32)));
        CK(h, cudaMalloc((void**)&h->json_uid_remap, kJsonUidSlots * sizeof(u32)));
        h->json_uid_dirty = true;
    }
    if (h->json_uid_dirty) {                                       // (12 MB: only after a log that had such uids)
        CK(h, cudaMemsetAsync(h->json_uid_keys, 0, kJsonUidSlots * sizeof(u64), st));
        CK(h, cudaMemsetAsync(h->json_uid_off, 0xFF, kJsonUidSlots * sizeof(u32), st));
        h->json_uid_dirty = false;
    }
    CK(h, cudaMemsetAsync(h->json_ns_keys, 0, kJsonNsSlots * sizeof(u64), st));
    CK(h, cudaMemsetAsync(h->json_ns_off, 0xFF, kJsonNsSlots * sizeof(u32), st));
    CK(h, cudaMemsetAsync(h->json_ph_keys, 0, kJsonPhSlots * sizeof(u64), st));
    CK(h, cudaMemsetAsync(h->json_ph_off, 0xFF, kJsonPhSlots * sizeof(u32), st));
    const u32 err_init[4] = {0u, 0xFFFFFFFFu, 0u, 0u};        // [2] is the parse kernel's line-group ticket
    CK(h, cudaMemcpyAsync(h->json_err, err_init, sizeof(err_init), cudaMemcpyHostToDevice, st));
    JsonArgs a;
    a.text = h->json_text; a.len = len; a.nl = h->json_nl; a.n_lines = n_lines; a.n_nl = n_nl;
    a.uid_hi = h->uid_hi; a.uid_lo = h->uid_lo; a.rv = h->rv; a.ns = h->ns; a.phase = h->phase; a.type = h->type;
    a.ns_dict = StrDict{h->json_ns_keys, h->json_ns_off, kJsonNsSlots - 1};
    a.phase_dict = StrDict{h->json_ph_keys, h->json_ph_off, kJsonPhSlots - 1};
    a.uid_dict = StrDict{h->json_uid_keys, h->json_uid_off, kJsonUidSlots - 1};
    a.err = h->json_err;
    a.arrival_order = h->json_stream ? 1 : 0; a.arrival_base = h->json_seq;
    grid = (int)std::min<u64>((n_lines + kJsonThreads - 1) / kJsonThreads, (u64)h->num_sms * 8);
    CK(h, cudaEventRecord(h->ev[1], st));
    json_parse_kernel<<<grid, kJsonThreads, 0, st>>>(a, h->json_err + 2);
    CK(h, cudaGetLastError());
    CK(h, cudaEventRecord(h->ev[2], st));
    u32 err[4];                                                   // [3]: rows whose uid went through the uid dictionary
    CK(h, cudaMemcpyAsync(err, h->json_err, sizeof(err), cudaMemcpyDeviceToHost, st));
    std::vector<u64> nk(kJsonNsSlots), pk(kJsonPhSlots);
    std::vector<u32> no(kJsonNsSlots), po(kJsonPhSlots);
    CK(h, cudaMemcpyAsync(nk.data(), h->json_ns_keys, nk.size() * 8, cudaMemcpyDeviceToHost, st));
    CK(h, cudaMemcpyAsync(no.data(), h->json_ns_off, no.size() * 4, cudaMemcpyDeviceToHost, st));
    CK(h, cudaMemcpyAsync(pk.data(), h->json_ph_keys, pk.size() * 8, cudaMemcpyDeviceToHost, st));
    CK(h, cudaMemcpyAsync(po.data(), h->json_ph_off, po.size() * 4, cudaMemcpyDeviceToHost, st));
    CK(h, cudaStreamSynchronize(st));
    if (err[3]) h->json_uid_dirty = true;
    if (err[0]) {
        static const char* why[] = {"", "malformed JSON", "event type is not ADDED/MODIFIED/DELETED", "metadata.uid is missing or not a string",
                                    "metadata.resourceVersion is not a decimal number", "a field the path reads contains a broken escape sequence",
                                    "more distinct strings than a device dictionary holds (65 536 namespaces, 1 048 576 uids that are not canonical UUIDs per log)", "too many distinct phase strings"};
        if (bad_line) *bad_line = err[1];
        h->n_events = 0;
        FAIL(h, PTE_ERR_PARSE, std::string("watch log line ") + std::to_string(err[1]) + ": " + why[std::min<u32>(err[0], 7)]);
    }
    // 3. dictionary slots -> dense ids in order of first occurrence (deterministic), names kept for the host
    auto build = [&](const std::vector<u64>& keys, const std::vector<u32>& offs, std::vector<std::string>& names, std::vector<u32>& remap,
                     std::unordered_map<std::string, u32>& ids) {
        std::vector<std::pair<u32, u32>> occ;      // (first offset, slot)
        for (u32 sl = 0; sl < keys.size(); sl++) if (keys[sl]) occ.emplace_back(offs[sl], sl);
        std::sort(occ.begin(), occ.end());
        remap.assign(keys.size(), 0);
        for (u32 id = 0; id < occ.size(); id++) {
            const u32 sl = occ[id].second, l = (u32)(keys[sl] & 0xFFFFFu);
            std::string name;                     // l DECODED bytes: the first occurrence may be spelled with escapes
            const HostJsonText ht{text};
            ptejson::Unescape<HostJsonText> u(ht, occ[id].first);
            for (u32 k = 0; k < l; k++) name.push_back((char)u.next());
            if (!h->json_stream) { names.push_back(std::move(name)); remap[sl] = id; continue; }
            // stream: a name keeps the id it got when the stream first saw it
            auto it = ids.find(name);
            if (it == ids.end()) { it = ids.emplace(name, (u32)names.size()).first; names.push_back(std::move(name)); }
            remap[sl] = it->second;
        }
    };
    std::vector<u32> nremap, premap32;
    build(nk, no, h->json_names[0], nremap, h->json_ids[0]);
    build(pk, po, h->json_names[1], premap32, h->json_ids[1]);
    if (h->json_names[1].size() > kPhMax - 6) {
        h->n_events = 0;
        FAIL(h, PTE_ERR_PARSE, "more than 56 distinct phase strings besides the five well-known ones");
    }
    std::vector<u8> premap(kJsonPhSlots, 0);
    for (u32 sl = 0; sl < kJsonPhSlots; sl++) premap[sl] = (u8)(7 + premap32[sl]);
    CK(h, cudaMemcpyAsync(h->json_ns_remap, nremap.data(), nremap.size() * 4, cudaMemcpyHostToDevice, st));
    CK(h, cudaMemcpyAsync(h->json_ph_remap, premap.data(), premap.size(), cudaMemcpyHostToDevice, st));
    const u32* uid_remap = nullptr;
    if (err[3]) {
        // some uids are not canonical UUIDs: their dictionary slots -> keys of the reserved block (1, 2, ... in order of first
        // occurrence; in a stream a string keeps the key it got first)
        std::vector<u64> uk(kJsonUidSlots);
        std::vector<u32> uo(kJsonUidSlots), uremap;
        CK(h, cudaMemcpyAsync(uk.data(), h->json_uid_keys, uk.size() * 8, cudaMemcpyDeviceToHost, st));
        CK(h, cudaMemcpyAsync(uo.data(), h->json_uid_off, uo.size() * 4, cudaMemcpyDeviceToHost, st));
        CK(h, cudaStreamSynchronize(st));
        build(uk, uo, h->json_names[2], uremap, h->json_ids[2]);
        CK(h, cudaMemcpyAsync(h->json_uid_remap, uremap.data(), uremap.size() * 4, cudaMemcpyHostToDevice, st));
        uid_remap = h->json_uid_remap;
    }
    grid = (int)std::min<u64>((n_lines + 255) / 256, (u64)h->num_sms * 16);
    json_remap_kernel<<<grid, 256, 0, st>>>(h->ns, h->phase, n_lines, h->json_ns_remap, h->json_ph_remap, h->uid_hi, h->uid_lo, uid_remap);
    CK(h, cudaGetLastError());
    u64 ends[2] = {0, 0};
    CK(h, cudaMemcpyAsync(&ends[0], h->rv, sizeof(u64), cudaMemcpyDeviceToHost, st));
    CK(h, cudaMemcpyAsync(&ends[1], h->rv + n_lines - 1, sizeof(u64), cudaMemcpyDeviceToHost, st));
    CK(h, cudaStreamSynchronize(st));
    h->n_events = n_lines;
    h->batch_first_rv = ends[0]; h->batch_last_rv = ends[1];
    cudaEventElapsedTime(&h->ms_json[0], h->ev[0], h->ev[1]);       // line index (count, scan, list) + dictionary set-up
    cudaEventElapsedTime(&h->ms_json[1], h->ev[1], h->ev[2]);       // json_parse_kernel
    h->json_bytes = len;
    if (h->json_stream) h->json_seq += n_lines;
    if (n_events) *n_events = n_lines;
    return PTE_OK;
}

int pte_json_stream(pte_handle h, int on) {
    if (!h) FAIL(h, PTE_ERR_INVALID, "handle is NULL");
    h->json_stream = on != 0;
    h->json_seq = 0;
    h->carry_valid = false;          // a stream carried so far (PTE_RUN_CARRY) ends here: its order keys mean nothing to the next one
    for (int k = 0; k < 3; k++) { h->json_names[k].clear(); h->json_ids[k].clear(); }
    return PTE_OK;
}

int pte_payload_timing(pte_handle h, float* ms_write, uint64_t* body_bytes) {
    if (!h) FAIL(h, PTE_ERR_INVALID, "handle is NULL");
    if (ms_write) *ms_write = h->ms_payload;
    if (body_bytes) *body_bytes = h->payload_bytes;
    return PTE_OK;
}

int pte_json_timing(pte_handle h, float* ms_index, float* ms_parse, uint64_t* text_bytes) {
    if (!h) FAIL(h, PTE_ERR_INVALID, "handle is NULL");
    if (ms_index) *ms_index = h->ms_json[0];
    if (ms_parse) *ms_parse = h->ms_json[1];
    if (text_bytes) *text_bytes = h->json_bytes;
    return PTE_OK;
}

int pte_json_strings(pte_handle h, int which, char* buf, uint64_t cap, uint32_t* offsets, uint32_t max_entries, uint32_t* n_out) {
    if (!h || !n_out || which < 0 || which > 2) FAIL(h, PTE_ERR_INVALID, "bad argument");
    const auto& v = h->json_names[which];
    *n_out = (u32)v.size();
    if (!buf || !offsets) return PTE_OK;
    if (v.size() > max_entries) FAIL(h, PTE_ERR_CAPACITY, "offset array too small");
    u64 o = 0;
    for (size_t i = 0; i < v.size(); i++) {
        if (o + v[i].size() > cap) FAIL(h, PTE_ERR_CAPACITY, "string buffer too small");
        offsets[i] = (u32)o;
        memcpy(buf + o, v[i].data(), v[i].size());
        o += v[i].size();
    }
    offsets[v.size()] = (u32)o;
    return PTE_OK;
}

void* pte_host_alloc(size_t bytes) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) {
        g_err = "cudaHostAlloc failed";
        cudaGetLastError();
        return nullptr;
    }
    return p;
}
void pte_host_free(void* p) { if (p) cudaFreeHost(p); }

uint32_t pte_shard_of(uint64_t uid_hi, uint64_t uid_lo, uint32_t n_shards) {
    (void)uid_lo;
    return n_shards ? shard_of_uid(uid_hi, n_shards) : 0;
}

int pte_debug_sort_u64(pte_handle h, uint64_t* records, uint64_t n, int bit_lo, int bit_hi) {
    if (!h) FAIL(h, PTE_ERR_INVALID, "handle is NULL");
    if (n > h->max_events) FAIL(h, PTE_ERR_CAPACITY, "more records than the engine's max_events");
    if (bit_lo < 0 || bit_hi > 64 || bit_hi < bit_lo) FAIL(h, PTE_ERR_INVALID, "bad bit range");
    if (n == 0) return PTE_OK;
    CK(h, cudaSetDevice(h->device));
    CK(h, cudaMemcpyAsync(h->recA, records, n * 8, cudaMemcpyHostToDevice, h->stream));
    u64* sorted = nullptr;
    h->pass_ev_used = 0;
    int rc = sort_field(h, h->recA, h->recB, h->recC, nullptr, n, bit_lo, bit_hi, &sorted);
    if (rc) return rc;
    CK(h, cudaMemcpyAsync(records, sorted, n * 8, cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    h->has_run = false;
    return PTE_OK;
}

int pte_comm_unique_id(char id_out[128]) {
    if (!id_out) FAIL((pte_handle) nullptr, PTE_ERR_INVALID, "id_out is NULL");
    if (!g_nccl.load(g_err)) return PTE_ERR_NCCL;
    ncclUniqueId id;
    ncclResult_t r = g_nccl.GetUniqueId(&id);
    if (r != ncclSuccess) FAIL((pte_handle) nullptr, PTE_ERR_NCCL, std::string("ncclGetUniqueId: ") + g_nccl.GetErrorString(r));
    static_assert(sizeof(id) == 128, "ncclUniqueId is 128 bytes");
    memcpy(id_out, &id, 128);
    return PTE_OK;
}

int pte_comm_init(pte_handle h, const char id[128], int rank, int n_ranks) {
    if (!h || !id) FAIL(h, PTE_ERR_INVALID, "NULL argument");
    if (n_ranks < 1 || rank < 0 || rank >= n_ranks) FAIL(h, PTE_ERR_INVALID, "rank must be in [0, n_ranks)");
    if (!g_nccl.load(h->err)) return PTE_ERR_NCCL;
    CK(h, cudaSetDevice(h->device));
    ncclUniqueId uid;
    memcpy(&uid, id, 128);
    ncclResult_t r = g_nccl.CommInitRank(&h->comm, n_ranks, uid, rank);
    if (r != ncclSuccess) FAIL(h, PTE_ERR_NCCL, std::string("ncclCommInitRank: ") + g_nccl.GetErrorString(r));
    h->rank = rank; h->n_ranks = n_ranks;
    if (h->d_counts_recv) { cudaFree(h->d_counts_recv); h->d_counts_recv = nullptr; }
    CK(h, cudaMalloc((void**)&h->d_counts_recv, sizeof(u64) * 4 * n_ranks));
    return PTE_OK;
}

int pte_gather_counts(pte_handle h, uint64_t* counts) {
    if (!h || !counts) FAIL(h, PTE_ERR_INVALID, "NULL argument");
    if (!h->has_run) FAIL(h, PTE_ERR_STATE, "pte_gather_counts before a successful pte_run");
    u64 mine[4] = {h->n_events, h->last.n_survived, h->last.n_emitted, h->last.n_tombstones};
    if (h->n_ranks == 1 && !h->comm) { memcpy(counts, mine, sizeof(mine)); return PTE_OK; }
    if (!h->comm) FAIL(h, PTE_ERR_STATE, "pte_comm_init has not been called");
    CK(h, cudaSetDevice(h->device));
    CK(h, cudaMemcpyAsync(h->d_counts_send, mine, sizeof(mine), cudaMemcpyHostToDevice, h->stream));
    ncclResult_t r = g_nccl.AllGather(h->d_counts_send, h->d_counts_recv, 4, ncclUint64, h->comm, h->stream);
    if (r != ncclSuccess) FAIL(h, PTE_ERR_NCCL, std::string("ncclAllGather: ") + g_nccl.GetErrorString(r));
    CK(h, cudaMemcpyAsync(counts, h->d_counts_recv, sizeof(u64) * 4 * h->n_ranks, cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    return PTE_OK;
}

}  // extern "C"
