/*
 * pte.h -- C ABI of the B200 bulk pod phase-transition engine (libpte.so).
 *
 * The reference (highreso-gpu/k8s-watcher) is pure Python and has no FFI or
 * plugin interface; its per-event path is PodWatcher.handle_pod_event
 * (watcher/pod_watcher.py:214-241).  This header is the boundary a maintainer
 * would bind with ctypes to run that path in bulk on a GPU.  Each entry point
 * cites the reference code it stands in for.
 *
 * Conventions: every call returns 0 (PTE_OK) or a negative pte_status; the
 * text for the last failure on a handle is pte_last_error(h) (h == NULL: the
 * last failure of a call that had no handle, per thread).  No exceptions cross
 * the boundary.  The caller owns all host buffers; the engine owns all device
 * memory for the lifetime of the handle.  A handle is bound to one device and
 * is not thread-safe: one host thread per handle.  There is NO CPU fallback:
 * without a CUDA device pte_create fails with PTE_ERR_NO_DEVICE.
 *
 * Replay-log record (one pod-watch event; the fields handle_pod_event reads,
 * pod_watcher.py:162-167,216-217, plus resourceVersion as the order key):
 *   uid_hi,uid_lo  u64,u64  metadata.uid as a 128-bit integer (never 0:0)
 *   rv             u64      metadata.resourceVersion; replay order is ascending rv, ties by arrival
 *                           (a log that is not in that order is pre-sorted on the device)
 *   ns_id          u32      metadata.namespace, interned by the host log
 *   phase          u8       PTE_PH_* code (0..62)
 *   type           u8       PTE_T_ADDED / MODIFIED / DELETED (event['type'], :267)
 */
#ifndef PTE_H
#define PTE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct pte_engine* pte_handle;

enum pte_status {
    PTE_OK = 0,
    PTE_ERR_INVALID = -1,     /* bad argument */
    PTE_ERR_CUDA = -2,        /* CUDA runtime error (text in pte_last_error) */
    PTE_ERR_NO_DEVICE = -3,   /* no usable CUDA device: the engine has no CPU path */
    PTE_ERR_CAPACITY = -4,    /* more events than max_events / record layout does not fit 64 bits */
    PTE_ERR_TABLE_FULL = -5,  /* more distinct uids than max_pods */
    PTE_ERR_BAD_EVENT = -6,   /* type not in {0,1,2}, phase > 62 or uid == 0:0 (the reference would crash
                                 on such an object at pod_watcher.py:216) */
    PTE_ERR_UNORDERED = -7,   /* rv decreases along arrival order and cannot be pre-sorted: span too wide
                                 (bits(rv_max - rv_min) + bits(max_events) > 64), or inside a stream batch */
    PTE_ERR_STATE = -8,       /* call out of sequence (e.g. fetch before run) */
    PTE_ERR_NCCL = -9,
    PTE_ERR_PARSE = -10       /* a line of a raw watch log could not be turned into a replay-log row */
};

enum { PTE_T_ADDED = 0, PTE_T_MODIFIED = 1, PTE_T_DELETED = 2 };
enum {
    PTE_PH_PENDING = 0, PTE_PH_RUNNING = 1, PTE_PH_SUCCEEDED = 2, PTE_PH_FAILED = 3, PTE_PH_UNKNOWN = 4,
    PTE_PH_NULL = 5,      /* status present, status.phase is None */
    PTE_PH_NOSTATUS = 6,  /* pod.status is None: payload phase "Unknown" (pod_watcher.py:167) */
    PTE_PH_FIRST_DYNAMIC = 7, PTE_PH_MAX = 62
};

typedef struct pte_stats {
    uint64_t n_in;          /* events in the log */
    uint64_t n_survived;    /* passed should_process_event (:204-212) and the namespace list (:226-229) */
    uint64_t n_emitted;     /* notifications (R-trace if dedupe == 0, T-trace otherwise) */
    uint64_t n_tombstones;  /* emitted DELETED events */
    uint64_t n_pods;        /* distinct uids in the dictionary (cumulative over a PTE_RUN_CARRY stream) */
    uint32_t ordered;       /* 1: rv was non-decreasing in arrival order; 0: the log was pre-sorted by rv */
    uint32_t n_launches;    /* kernels launched by this pte_run */
    float ms_total;         /* device time of this pte_run, CUDA events on the engine stream */
    float ms_intern;        /* filter + uid interning + record packing */
    float ms_sort_key;      /* LSD radix passes on the interned key */
    float ms_diff;          /* segmented adjacent-diff -> emit bitmap (dedupe == 0: the predicate bitmap kernel) */
    float ms_sort_order;    /* always 0 since replay order is restored by the bitmap scan in ms_assemble */
    float ms_assemble;      /* bitmap scan + compaction in replay order + payload columns */
    float ms_pass_avg;      /* mean duration of one radix scatter pass (the dominant kernel) */
    uint32_t n_passes;      /* radix scatter passes run */
    uint32_t n_diff_rounds; /* diff_flag_kernel launches (one per bitmap slice; one for two slices when the second is deferred) */
    uint32_t path;          /* 0: sort path (intern, radix sort, adjacent diff); 1: partition-first path (large one-shot T-trace runs).
                               On path 1 ms_intern = partition counts + scatter, ms_sort_key = 0, ms_diff = replay of the partitions */
    uint32_t n_pending;     /* path 1: events resolved by the cross-segment fix-up (cut partitions only) */
    float ms_part_hist;     /* path 1: uid hash -> partition counts + replay-order check + segment table */
    float ms_part_scatter;  /* path 1: stable scatter of the events into 24-byte partition records */
    float ms_part_replay;   /* path 1: in-order replay of every partition against shared-memory state (+ fix-up) */
    uint32_t n_rebuilds;    /* PTE_RUN_CARRY streams: times the dictionary was rebuilt from its live pods so far */
    float ms_rebuild;       /* device time of the last such rebuild */
} pte_stats;

/* Parameters of the synthetic replay log (see "Synthetic replay logs" below). */
typedef struct pte_gen_params {
    uint64_t seed;
    uint64_t n_events;       /* events generated for this shard */
    uint64_t n_pods;         /* pods of this shard (pod index q in [0, n_pods)) */
    uint32_t dist;           /* 0 uniform pod activity, 1 octave-Zipf (P(rank r) ~ 1/(r+1)) */
    uint32_t p_del_q24;      /* delete-draw probability * 2^24 */
    uint32_t p_nostatus_q24; /* P(status is None) * 2^24 */
    uint32_t p_null_q24;     /* P(status.phase is None) * 2^24 */
    uint32_t shard;          /* this shard, in [0, n_shards) */
    uint32_t n_shards;       /* uid-hash shards (one per GPU) */
} pte_gen_params;

/* ---- lifetime ---------------------------------------------------------- */

/* Number of CUDA devices (0 if none / no driver). */
int pte_device_count(void);

/* Engine for up to max_events events per log and max_pods distinct uids, on `device`.
 * Stands in for PodWatcher.__init__ (pod_watcher.py:11-17) as far as the path needs state. */
int pte_create(int device, uint64_t max_events, uint64_t max_pods, pte_handle* out);
void pte_destroy(pte_handle h);
const char* pte_last_error(pte_handle h);

/* ---- filters ------------------------------------------------------------ */

/* The two per-event predicates of the reference, as data:
 *   production_critical_only = (environment == 'production' && watcher.alerts.critical_events_only)
 *                              -> should_process_event, pod_watcher.py:204-212
 *   ns_allow[0..n_allow)     = ids of watcher.namespaces; n_allow == 0 means "all namespaces"
 *                              -> pod_watcher.py:226-229
 *   ns_filter_on             = 1 if the reference's list was non-empty even though none of its
 *                              names occurs in the log (then every event is dropped, as upstream). */
int pte_set_filters(pte_handle h, const uint32_t* ns_allow, int n_allow, int ns_filter_on,
                    int production_critical_only);

/* ---- input -------------------------------------------------------------- */

/* Copy a log of n events from host columns to the device (replaces the current log).
 * Stands in for the `for event in stream` producer, pod_watcher.py:264-268.
 * ns_id may be NULL while no namespace allow-list is set (nothing on the path reads it then,
 * pod_watcher.py:226-229 is a no-op; the ns_id output column reads 0): saves 4 of the 30 bytes per event. */
int pte_ingest(pte_handle h, const uint64_t* uid_hi, const uint64_t* uid_lo, const uint64_t* rv,
               const uint32_t* ns_id, const uint8_t* phase, const uint8_t* type, uint64_t n);

/* Parse a raw watch log on the device: newline-delimited JSON watch events, one per line, as the Kubernetes
 * API streams them and the reference's loop receives them (event['type'], event['object'], pod_watcher.py:264-268):
 *   {"type":"MODIFIED","object":{"metadata":{"uid":..,"namespace":..,"resourceVersion":..},"status":{"phase":..},..}}
 * Keys in any order, any whitespace; the rest of the object is skipped (nesting- and string-aware).  Replaces
 * the current log.  metadata.uid in the form Kubernetes emits (lower-case hex, 8-4-4-4-12) is its own 128-bit key; any other
 * uid string -- the reference takes whatever arrives, pod_watcher.py:164 -- is interned exactly on the device and gets the key
 * (2^64 - 1, k), k = 1, 2, ... in order of first occurrence (pte_json_strings(which = 2) returns those strings, index k - 1;
 * canonical strings whose value lies in that block, and the all-zero uid, go the same way): two different uid strings never
 * share a key.  A missing resourceVersion becomes line number + 1; status absent or null -> PTE_PH_NOSTATUS, phase absent or null -> PTE_PH_NULL.  Namespaces, and phase
 * strings other than the five well-known ones, are interned exactly on the device (ids / codes >= 7 in order of
 * first occurrence); pte_json_strings returns the dictionaries.  JSON escape sequences in the five strings are decoded
 * (\" \\ \/ \b \f \n \r \t, \uXXXX with surrogate pairs, as UTF-8): two spellings of one string are one string, and
 * pte_json_strings returns the decoded bytes.  On a bad line (not one JSON container, type / resourceVersion not
 * of the form above, no uid string, a broken escape): PTE_ERR_PARSE, *bad_line = its 0-based number, and the log is left empty. */
int pte_ingest_json(pte_handle h, const char* text, uint64_t len, uint64_t* n_events, uint64_t* bad_line);

/* Raw watch text as a STREAM (the body of `for event in w.stream(...)`, pod_watcher.py:264-269, fed from the wire in pieces):
 * with on != 0 every following pte_ingest_json is the next batch of one stream -- rows are numbered by arrival across the
 * batches (that number is the order key; metadata.resourceVersion is not read, as in the reference's loop), and a
 * namespace, phase or non-UUID uid string keeps the id / code / key it got when the stream first showed it, so PTE_RUN_CARRY compares like with
 * like and pte_json_strings returns the stream's dictionaries so far.  Each call (on or off) starts afresh: names, numbering,
 * and the per-pod state carried by earlier PTE_RUN_CARRY runs. */
int pte_json_stream(pte_handle h, int on);

/* Device time of the last pte_ingest_json (CUDA events on the engine stream, host <-> device copies excluded):
 * the line index (count, scan, list of newlines) and the scanner kernel; text_bytes = size of that log. */
int pte_json_timing(pte_handle h, float* ms_index, float* ms_parse, uint64_t* text_bytes);

/* Dictionary of the last pte_ingest_json: which = 0 namespaces (index = ns_id), 1 phase strings (index = code - 7),
 * 2 uid strings that are not canonical UUIDs (index = uid_lo - 1 of the rows with uid_hi == 2^64 - 1).
 * Strings are written back to back into buf, offsets[0..n] delimit them; buf == NULL just counts (*n_out). */
int pte_json_strings(pte_handle h, int which, char* buf, uint64_t cap, uint32_t* offsets, uint32_t max_entries,
                     uint32_t* n_out);

/* Generate the synthetic log directly in device memory (no reference counterpart). */
int pte_generate(pte_handle h, const pte_gen_params* p);

/* The same generator on the host, strictly sequential, into caller buffers (no device needed). */
int pte_generate_host(const pte_gen_params* p, uint64_t* uid_hi, uint64_t* uid_lo, uint64_t* rv,
                      uint32_t* ns_id, uint8_t* phase, uint8_t* type);

/* Copy the device-resident log back (any pointer may be NULL); cap = capacity of each buffer. */
int pte_read_log(pte_handle h, uint64_t* uid_hi, uint64_t* uid_lo, uint64_t* rv, uint32_t* ns_id,
                 uint8_t* phase, uint8_t* type, uint64_t cap, uint64_t* n_out);

/* ---- the hot path --------------------------------------------------------- */

/* Replay the device-resident log: filters -> uid interning -> LSD radix sort on (key, rv) ->
 * segmented adjacent-diff -> emit bitmap -> scan of the bitmap in replay order -> payload columns.
 * flags = 0               : every event that would reach the POST site (pod_watcher.py:232-236), the R-trace.
 * flags & PTE_RUN_DEDUPE  : only phase transitions and DELETED tombstones (the last-known-phase layer the
 *                           reference's README describes but pod_watcher.py does not implement), the T-trace.
 * flags & PTE_RUN_CARRY   : (with DEDUPE) the log is the next batch of a stream: the uid dictionary and the
 *                           last known phase of every pod are kept from the previous PTE_RUN_CARRY call on
 *                           this handle, so that micro-batches of the live loop (pod_watcher.py:264-269) give
 *                           exactly the notifications one replay of the whole stream would.  A call without
 *                           the flag, or a failed call, ends the stream; the next CARRY call starts a new one.
 * Without PTE_RUN_CARRY the dictionary is cleared at the start of the call.
 * flags & PTE_RUN_BITMAP_ONLY : stop at the emit bitmap and the counts; the payload columns are not assembled
 *                           (pte_fetch, pte_digest and pte_serialize_payloads then answer PTE_ERR_STATE).  For hosts that
 *                           still hold the log and fetch the result with pte_fetch_bitmap. */
enum { PTE_RUN_DEDUPE = 1, PTE_RUN_CARRY = 2, PTE_RUN_BITMAP_ONLY = 4 };
int pte_run(pte_handle h, int flags, pte_stats* out);

/* Emitted records of the last pte_run, in replay order, into caller columns (any may be NULL):
 * idx = arrival index in the ingested log, key = interned 64-bit uid key (dedupe runs only; ~0 otherwise).
 * Stands in for the payload hand-off at pod_watcher.py:232-236. */
int pte_fetch(pte_handle h, uint32_t* idx, uint64_t* rv, uint64_t* key, uint32_t* ns_id,
              uint8_t* phase, uint8_t* type, uint64_t cap, uint64_t* n_out);

/* The same result as one bit per event: bit (i & 31) of words[i >> 5] is set iff event i of the resident log
 * reaches the notification site pod_watcher.py:232-236 (and, with dedupe, is a transition).  A host that still
 * holds the log needs nothing else for the payload hand-off, and downloads 1 bit per event instead of the
 * 5 bytes per notification of pte_fetch(idx, type).  words: (n_events + 31) / 32 uint32 (NULL: only *n_emitted).
 * Not available (PTE_ERR_STATE) when the run had to pre-sort the log by rv: use pte_fetch then. */
int pte_fetch_bitmap(pte_handle h, uint32_t* words, uint64_t cap_words, uint64_t* n_emitted);

/* ingest + run + fetch_bitmap in one call, pipelined: the host columns (pinned memory for full speed) go to the device
 * in chunks of chunk_events events (0: 2^26) on a copy stream of their own, and every chunk is replayed as the next
 * batch of one stream (PTE_RUN_CARRY) as soon as it has landed, so the copies of the later chunks hide behind the
 * replay of the earlier ones and the bitmap of every chunk is on its way back while the next is replayed.  Same
 * notifications as pte_ingest + pte_run on the whole log (dedupe as in pte_run); the emit bitmap lands in
 * bitmap_out ((n + 31) / 32 words, layout of pte_fetch_bitmap), the counts in *out (times are sums over chunks).
 * This is the payload hand-off of pod_watcher.py:232-236 for a host that holds the log.  The log must be in
 * resourceVersion order (PTE_ERR_UNORDERED otherwise); ns_id may be NULL as in pte_ingest.  rv may be NULL too: the
 * caller then vouches that the log is in arrival order -- the only order the reference's loop knows, it never reads
 * metadata.resourceVersion -- and the position serves as order key (8 of the 26 bytes per event stay at home).
 * type may be NULL as well: phase[] then carries both, type << 6 | phase in one byte per event (PTE_PACK_META below), and is
 * taken apart on the device -- 17 bytes per event over the wire.  The log stays resident. */
#define PTE_PACK_META(type, phase) ((uint8_t)(((type) << 6) | ((phase) & 63)))
int pte_replay_pipelined(pte_handle h, const uint64_t* uid_hi, const uint64_t* uid_lo, const uint64_t* rv,
                         const uint32_t* ns_id, const uint8_t* phase, const uint8_t* type, uint64_t n,
                         uint64_t chunk_events, int dedupe, uint32_t* bitmap_out, pte_stats* out);

/* One batch of a live stream (the body of `for event in stream`, pod_watcher.py:266-269, for a micro-batch) in one call:
 * pte_ingest + pte_run(flags, normally PTE_RUN_DEDUPE | PTE_RUN_CARRY) + pte_fetch(idx, type), with one host
 * synchronisation instead of three.  idx_out / type_out (capacity cap) receive the arrival index and the event type of
 * every notification of the batch; *out the counts.  Afterwards the engine is in the state pte_run leaves it in.
 * Batches of up to 4096 events are replayed by a single kernel launch (packed pinned staging in, one packed copy out;
 * same results; PTE_TINY=0 in the environment of pte_create switches that path off); pte_stats then holds ms_total and
 * n_launches = 1 only.  Stage times (ms_intern ...) are measured for runs of 2^20 events or more (PTE_STAGE_TIMING=1: always). */
int pte_stream_batch(pte_handle h, const uint64_t* uid_hi, const uint64_t* uid_lo, const uint64_t* rv,
                     const uint32_t* ns_id, const uint8_t* phase, const uint8_t* type, uint64_t n, int flags,
                     uint32_t* idx_out, uint8_t* type_out, uint64_t cap, pte_stats* out);

/* Raise max_events of a live engine without touching the uid dictionary or the per-pod state a PTE_RUN_CARRY
 * stream carries: the live loop (pod_watcher.py:264-269) may meet a longer batch than any before.  The resident
 * log and the results of the last run are dropped; ingest again.  No-op when max_events is not larger. */
int pte_grow(pte_handle h, uint64_t max_events);

/* The request bodies the reference would POST for emitted records [first, first + count) of the last pte_run:
 * json.dumps(_extract_pod_data(pod) + event_type) (pod_watcher.py:159-202, :233; clusterapi_client.py:36),
 * for pods that carry only what a replay-log row holds (name "pod-" + uid[:13], no labels, containers, ...).
 * The JSON text around the per-record fields is given by the caller as ready-made fragments, so nothing has
 * to be escaped on the device: blob holds 6 segments, then 3 event-type fragments, n_phase phase fragments
 * (by PTE_PH_* code) and n_ns namespace fragments (by ns_id), back to back; fragment i is
 * blob[frag_off[i], frag_off[i+1]).  A body is  seg0 uid[:13] seg1 NS seg2 uid seg3 PHASE seg4 TYPE seg5.
 * out receives the bodies back to back, offsets[0..count] their start (offsets[count] = total bytes);
 * with out == NULL only *n_bytes (the size needed) is computed. */
int pte_serialize_payloads(pte_handle h, uint64_t first, uint64_t count, const char* blob, uint32_t blob_len,
                           const uint32_t* frag_off, uint32_t n_phase, uint32_t n_ns, char* out, uint64_t out_cap,
                           uint64_t* offsets, uint64_t* n_bytes);

/* Device time of the kernel that wrote the bodies of the last pte_serialize_payloads (CUDA events, copies excluded). */
int pte_payload_timing(pte_handle h, float* ms_write, uint64_t* body_bytes);

/* Order-sensitive 64-bit digest of the emitted records (position, rv, uid, ns, phase, type);
 * same arithmetic as oracle/oracle.c rec_digest.  For parity at sizes too big to compare in full. */
int pte_digest(pte_handle h, uint64_t* out);

/* ---- host helpers ------------------------------------------------------- */

/* Pinned host memory for ingest / fetch buffers (plain malloc'd buffers work too, slower). */
void* pte_host_alloc(size_t bytes);
void pte_host_free(void* p);

/* Shard of a uid when a log is split across n_shards engines (one per GPU). */
uint32_t pte_shard_of(uint64_t uid_hi, uint64_t uid_lo, uint32_t n_shards);

/* Stable LSD radix sort of n u64 records on bits [bit_lo, bit_hi), through the device sort
 * kernels, in place in the caller's host array.  Test hook for the sort in isolation. */
int pte_debug_sort_u64(pte_handle h, uint64_t* records, uint64_t n, int bit_lo, int bit_hi);

/* No counterpart in the reference.
 * One stable radix scatter pass that allocates nothing, for hosts where this library must not
 * call cudaMalloc: servers that give all device memory to their own pool at start-up, and plug-ins
 * on latency-critical paths that cannot take the device-wide stall an allocation may cause.
 * out receives the n records of `in` stably partitioned by the digit in bits [bit_lo, bit_lo + n_bits)
 * (n_bits 1..8) -- the same result as pte_debug_sort_u64 on that bit range.  in and out are distinct,
 * caller-owned, 8-byte aligned and reachable from the GPU (device memory, or pinned host memory such as
 * pte_host_alloc returns); no padding is needed.  The work runs on `stream` (a cudaStream_t; NULL is the
 * legacy default stream) of the current device; the call waits for that stream (cudaStreamSynchronize)
 * before it returns, so the result is in `out` when it does.
 * The histogram and look-back scratch is a fixed array inside the library's device image, about 1 MB that
 * every GPU which loads libpte.so holds until the process exits.  Hence:
 *   - n may not exceed PTE_STATIC_MAX_RECORDS; more is refused with PTE_ERR_CAPACITY;
 *   - all calls on one GPU share the array (each call first clears what it will read), so two calls
 *     that overlap on the same GPU -- from two threads or two streams -- race: serialise them;
 *   - the first call on each GPU looks up the array's device address and sets kernel attributes;
 *     make that call before any stream capture begins.
 * On failure the reason is copied to err (capacity err_cap) when err is not NULL. */
#define PTE_STATIC_MAX_RECORDS (1ull << 22)
int pte_partition_u64_static(const uint64_t* in, uint64_t* out, uint64_t n, int bit_lo, int n_bits,
                             void* stream, char* err, size_t err_cap);

/* ---- multi-GPU: one engine per process/GPU, a single collective --------- */

/* NCCL bootstrap: rank 0 makes an id, the host layer ships its 128 bytes to every rank. */
int pte_comm_unique_id(char id_out[128]);
int pte_comm_init(pte_handle h, const char id[128], int rank, int n_ranks);
/* All-gather of {n_in, n_survived, n_emitted, n_tombstones} of the last pte_run:
 * counts[4 * n_ranks], rank-major.  The only data that crosses GPUs on this path. */
int pte_gather_counts(pte_handle h, uint64_t* counts);

/*
 * Synthetic replay logs (shared spec of pte_generate, pte_generate_host and oracle/oracle.c).
 * All integer; mix64 is the splitmix64 finaliser.  For local event j of shard g of G:
 *   e   = j*G + g                      rv = e + 1
 *   r1  = mix64(seed ^ mix64(e + 0x9E3779B97F4A7C15)), r2 = mix64(r1 + 0xD1B54A32D192ED03),
 *   r3  = mix64(r2 + 0x8CB92BA72F3D8DD7),              r4 = mix64(r3 + 0x2545F4914F6CDD1D)
 *   q   = mulhi64(r1, n_pods)                                        (dist 0)
 *       = ((2^o - 1 + (r2 & (2^o - 1))) mod n_pods * 2654435761 + 40503) mod n_pods,
 *         o = mulhi64(r1, bitlen(n_pods))                            (dist 1)
 *   pod = q*G + g    uid_hi = mix64(pod ^ 0x6B38735F77617463)  uid_lo = mix64(~pod ^ 0x5F62323030706F64)
 *   ns_id = pod mod 5    phase = ((r3 & 0xFFFFFFFF) * 5) >> 32, overridden by NOSTATUS if
 *   (r4 & 0xFFFFFF) < p_nostatus_q24, else by NULL if ((r4 >> 24) & 0xFFFFFF) < p_null_q24
 *   draw = (r3 >> 40) < p_del_q24
 *   type = ADDED if the pod has no earlier event or its previous event had draw set,
 *          else DELETED if draw, else MODIFIED.
 */

#ifdef __cplusplus
}
#endif
#endif /* PTE_H */
