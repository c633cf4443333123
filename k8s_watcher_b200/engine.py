"""Python face of the C ABI: one :class:`PhaseTransitionEngine` per GPU.

Everything here is argument marshalling; the work happens in libpte.so
(k8s_watcher_b200/csrc).  Without the library or without a CUDA device the
constructor raises -- there is no CPU path.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Dict, Iterable, Optional, Sequence

import numpy as np

from . import _abi
from ._abi import PteError, PteGenParams, PteStats  # noqa: F401  (re-exported)

COLUMNS = (("uid_hi", np.uint64), ("uid_lo", np.uint64), ("rv", np.uint64),
           ("ns_id", np.uint32), ("phase", np.uint8), ("type", np.uint8))


def q24(p: float) -> int:
    return int(round(float(p) * (1 << 24)))


def gen_params(seed: int, n_events: int, n_pods: int, *, dist: int = 0, p_del: float = 0.0, p_nostatus: float = 0.0,
               p_null: float = 0.0, shard: int = 0, n_shards: int = 1) -> PteGenParams:
    return PteGenParams(seed & 0xFFFFFFFFFFFFFFFF, n_events, n_pods, dist, q24(p_del), q24(p_nostatus), q24(p_null),
                        shard, n_shards)


def _ptr(a: Optional[np.ndarray], ctype):
    return None if a is None else a.ctypes.data_as(C.POINTER(ctype))


def pack_meta(type_col: np.ndarray, phase_col: np.ndarray, out: Optional[np.ndarray] = None) -> np.ndarray:
    """One byte per event for event type and phase: ``type << 6 | phase`` (PTE_PACK_META of include/pte.h)."""
    out = np.empty(type_col.shape[0], np.uint8) if out is None else out
    np.left_shift(type_col, 6, out=out, dtype=np.uint8, casting="unsafe")
    np.bitwise_or(out, phase_col & np.uint8(63), out=out)
    return out


def generate_host(seed: int, n_events: int, n_pods: int, **kw) -> Dict[str, np.ndarray]:
    """The synthetic replay log of include/pte.h, generated on the host by libpte.so."""
    cols = {name: np.empty(n_events, dt) for name, dt in COLUMNS}
    p = gen_params(seed, n_events, n_pods, **kw)
    rc = _abi.lib().pte_generate_host(C.byref(p), _ptr(cols["uid_hi"], C.c_uint64), _ptr(cols["uid_lo"], C.c_uint64),
                                      _ptr(cols["rv"], C.c_uint64), _ptr(cols["ns_id"], C.c_uint32),
                                      _ptr(cols["phase"], C.c_uint8), _ptr(cols["type"], C.c_uint8))
    _abi.check(rc)
    return cols


def shard_of(uid_hi: int, uid_lo: int, n_shards: int) -> int:
    return int(_abi.lib().pte_shard_of(int(uid_hi), int(uid_lo), int(n_shards)))


STATIC_MAX_RECORDS = 1 << 22


def partition_u64_static(src: np.ndarray, dst: np.ndarray, bit_lo: int, n_bits: int, stream: int = 0) -> None:
    """pte_partition_u64_static: one stable radix pass that allocates nothing on the device.

    For hosts that own all device memory themselves and forbid library allocations.  ``src`` and ``dst``
    must be GPU-reachable uint64 arrays of the same length (e.g. ``PinnedBuffer(...).array``); at most
    ``STATIC_MAX_RECORDS`` records; calls on one GPU must not overlap (they share one scratch array).
    """
    if src.dtype != np.uint64 or dst.dtype != np.uint64 or src.shape != dst.shape or src.ndim != 1:
        raise ValueError("src and dst must be 1-d uint64 arrays of equal length")
    err = C.create_string_buffer(256)
    rc = _abi.lib().pte_partition_u64_static(_ptr(src, C.c_uint64), _ptr(dst, C.c_uint64), int(src.shape[0]),
                                             int(bit_lo), int(n_bits), C.c_void_p(stream), err, 256)
    if rc != _abi.PTE_OK:
        raise PteError(rc, err.value.decode(errors="replace"))


class PinnedBuffer:
    """Page-locked host array (pte_host_alloc) viewed through numpy."""

    def __init__(self, count: int, dtype):
        self.dtype = np.dtype(dtype)
        self.nbytes = max(1, count * self.dtype.itemsize)
        self._ptr = _abi.lib().pte_host_alloc(self.nbytes)
        if not self._ptr:
            raise MemoryError("pte_host_alloc(%d) failed: %s" % (self.nbytes, _abi.last_error(None)))
        buf = (C.c_uint8 * self.nbytes).from_address(self._ptr)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=count)

    def free(self):
        if self._ptr:
            self.array = None
            _abi.lib().pte_host_free(self._ptr)
            self._ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


@dataclass
class EmittedColumns:
    """Emitted records in replay order (what the sink is handed)."""
    idx: np.ndarray      # arrival index in the ingested log
    rv: np.ndarray
    key: np.ndarray      # interned 64-bit uid key
    ns_id: np.ndarray
    phase: np.ndarray
    type: np.ndarray

    def __len__(self):
        return int(self.idx.shape[0])


class PhaseTransitionEngine:
    """One engine handle, bound to one GPU.  Not thread-safe."""

    def __init__(self, max_events: int, max_pods: int = 0, device: int = 0):
        L = _abi.lib()
        if _abi.device_count() <= 0:
            raise PteError(-3, "no CUDA device: the phase-transition engine has no CPU path")
        self._h = C.c_void_p()
        _abi.check(L.pte_create(int(device), int(max_events), int(max_pods), C.byref(self._h)))
        self._L = L
        self._stream_batch_fn = None
        self.max_events, self.max_pods, self.device = int(max_events), int(max_pods) or int(max_events), int(device)
        self.last_stats: Optional[Dict[str, float]] = None

    # -- lifetime
    def close(self):
        if getattr(self, "_h", None):
            self._L.pte_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc: int):
        _abi.check(rc, self._h)

    # -- configuration
    def set_filters(self, ns_allow: Optional[Iterable[int]] = None, *, ns_filter_on: bool = False,
                    critical_only: bool = False):
        ids = np.ascontiguousarray(sorted(set(int(x) for x in ns_allow)) if ns_allow else [], dtype=np.uint32)
        self._ck(self._L.pte_set_filters(self._h, _ptr(ids, C.c_uint32) if ids.size else None, int(ids.size),
                                         int(bool(ns_filter_on)), int(bool(critical_only))))

    # -- input
    def ingest(self, cols: Dict[str, np.ndarray], *, with_namespaces: bool = True):
        """Copy a log to the device.  ``with_namespaces=False`` leaves the ns_id column at home (allowed while no
        namespace allow-list is set): 26 instead of 30 bytes per event over PCIe."""
        n = int(cols["rv"].shape[0])
        c = {name: np.ascontiguousarray(cols[name], dtype=dt) for name, dt in COLUMNS}
        if not with_namespaces:
            c["ns_id"] = None
        for name, _ in COLUMNS:
            if c[name] is not None and c[name].shape[0] != n:
                raise ValueError("column %s has %d rows, rv has %d" % (name, c[name].shape[0], n))
        self._ck(self._L.pte_ingest(self._h, _ptr(c["uid_hi"], C.c_uint64), _ptr(c["uid_lo"], C.c_uint64),
                                    _ptr(c["rv"], C.c_uint64), _ptr(c["ns_id"], C.c_uint32),
                                    _ptr(c["phase"], C.c_uint8), _ptr(c["type"], C.c_uint8), n))

    def ingest_json(self, text) -> int:
        """Parse a raw watch log (newline-delimited JSON watch events) on the device into the log columns.
        ``text``: bytes, or a uint8 numpy array -- one in pinned memory (PinnedBuffer) reaches the device at full PCIe
        speed, which is what a gigabyte of text wants: the scanner itself takes 2 ms of it.
        Returns the number of events; raises PteError(PTE_ERR_PARSE) naming the first bad line."""
        n, bad = C.c_uint64(), C.c_uint64()
        if isinstance(text, np.ndarray):
            a = np.ascontiguousarray(text, dtype=np.uint8)
            self._ck(self._L.pte_ingest_json(self._h, C.cast(a.ctypes.data, C.c_char_p), a.nbytes, C.byref(n), C.byref(bad)))
        else:
            self._ck(self._L.pte_ingest_json(self._h, text, len(text), C.byref(n), C.byref(bad)))
        return int(n.value)

    def payload_timing(self) -> Dict[str, float]:
        """Device time of the last serialize_payloads' write kernel: {"ms_write", "body_bytes"} (pte_payload_timing)."""
        a, n = C.c_float(), C.c_uint64()
        self._ck(self._L.pte_payload_timing(self._h, C.byref(a), C.byref(n)))
        return {"ms_write": a.value, "body_bytes": int(n.value)}

    def json_stream(self, on: bool = True):
        """The following ingest_json calls are the batches of one stream: rows numbered by arrival across batches, ids of
        namespaces and codes of phase strings persistent (pte_json_stream)."""
        self._ck(self._L.pte_json_stream(self._h, 1 if on else 0))

    def json_timing(self) -> Dict[str, float]:
        """Device time of the last ingest_json: {"ms_index", "ms_parse", "text_bytes"} (pte_json_timing)."""
        a, b, n = C.c_float(), C.c_float(), C.c_uint64()
        self._ck(self._L.pte_json_timing(self._h, C.byref(a), C.byref(b), C.byref(n)))
        return {"ms_index": a.value, "ms_parse": b.value, "text_bytes": int(n.value)}

    def json_strings(self, which: int) -> list:
        """Dictionary of the last ingest_json: 0 = namespaces by ns_id, 1 = phase strings for codes 7, 8, ...,
        2 = uid strings that are not canonical UUIDs (index = uid_lo - 1 where uid_hi == 2**64 - 1)."""
        n = C.c_uint32()
        self._ck(self._L.pte_json_strings(self._h, which, None, 0, None, 0, C.byref(n)))
        if n.value == 0:
            return []
        off = np.zeros(n.value + 1, np.uint32)
        size = 1 << 20
        while True:                                   # the ABI has no size query for the text: grow until it fits
            buf = C.create_string_buffer(size)
            rc = self._L.pte_json_strings(self._h, which, buf, len(buf), _ptr(off, C.c_uint32), n.value, C.byref(n))
            if rc != -4 or size >= (1 << 32):         # -4 = PTE_ERR_CAPACITY
                break
            size <<= 2
        self._ck(rc)
        raw = buf.raw
        return [raw[int(off[i]):int(off[i + 1])].decode("utf-8") for i in range(n.value)]

    def generate(self, seed: int, n_events: int, n_pods: int, **kw):
        p = gen_params(seed, n_events, n_pods, **kw)
        self._ck(self._L.pte_generate(self._h, C.byref(p)))

    def read_log(self, into: Optional[Dict[str, np.ndarray]] = None) -> Dict[str, np.ndarray]:
        n = C.c_uint64()
        self._ck(self._L.pte_read_log(self._h, None, None, None, None, None, None, 1 << 62, C.byref(n)))
        cols = into or {name: np.empty(n.value, dt) for name, dt in COLUMNS}
        self._ck(self._L.pte_read_log(self._h, _ptr(cols["uid_hi"], C.c_uint64), _ptr(cols["uid_lo"], C.c_uint64),
                                      _ptr(cols["rv"], C.c_uint64), _ptr(cols["ns_id"], C.c_uint32),
                                      _ptr(cols["phase"], C.c_uint8), _ptr(cols["type"], C.c_uint8),
                                      int(cols["rv"].shape[0]), C.byref(n)))
        return cols

    # -- the hot path
    def run(self, dedupe: bool = True, carry: bool = False, bitmap_only: bool = False) -> Dict[str, float]:
        """Replay the resident log.  ``carry`` continues a stream: dictionary and per-pod last phase are kept
        from the previous ``carry`` run, so consecutive batches add up to one replay of the whole stream."""
        st = PteStats()
        flags = (1 if dedupe else 0) | (2 if carry and dedupe else 0) | (4 if bitmap_only else 0)
        self._ck(self._L.pte_run(self._h, flags, C.byref(st)))
        self.last_stats = st.as_dict()
        return self.last_stats

    ALL_COLUMNS = ("idx", "rv", "key", "ns_id", "phase", "type")

    def fetch(self, columns: Sequence[str] = ALL_COLUMNS,
              into: Optional[Dict[str, np.ndarray]] = None) -> EmittedColumns:
        """Emitted records of the last run.  ``idx`` (arrival index) identifies a record completely for a
        host that still holds the log; the other columns are there for sinks that do not."""
        n = int(self.last_stats["n_emitted"]) if self.last_stats else 0
        dts = dict(idx=np.uint32, rv=np.uint64, key=np.uint64, ns_id=np.uint32, phase=np.uint8, type=np.uint8)
        out = {}
        for name in dts:
            if into is not None and name in into:
                out[name] = into[name]
            elif name in columns:
                out[name] = np.empty(n, dts[name])
            else:
                out[name] = None
        cap = min(int(a.shape[0]) for a in out.values() if a is not None) if any(a is not None for a in out.values()) else n
        got = C.c_uint64()
        self._ck(self._L.pte_fetch(self._h, _ptr(out["idx"], C.c_uint32), _ptr(out["rv"], C.c_uint64),
                                   _ptr(out["key"], C.c_uint64), _ptr(out["ns_id"], C.c_uint32),
                                   _ptr(out["phase"], C.c_uint8), _ptr(out["type"], C.c_uint8), cap, C.byref(got)))
        k = int(got.value)
        z = {name: (np.empty(0, dts[name]) if a is None else a[:k]) for name, a in out.items()}
        return EmittedColumns(**z)

    def serialize_payloads(self, fragments: Sequence[bytes], n_phase: int, n_ns: int, first: int = 0,
                           count: Optional[int] = None):
        """Request bodies of emitted records [first, first+count) built on the device (pte_serialize_payloads).
        ``fragments`` = 6 segments + 3 event types + n_phase phases + n_ns namespaces, each ready-made JSON text.
        Returns (bytes buffer as uint8 array, offsets uint64[count+1])."""
        n = int(self.last_stats["n_emitted"]) if self.last_stats else 0
        count = n - first if count is None else count
        assert len(fragments) == 9 + n_phase + n_ns
        off = np.zeros(len(fragments) + 1, np.uint32)
        off[1:] = np.cumsum([len(f) for f in fragments])
        blob = b"".join(fragments)
        need = C.c_uint64()
        self._ck(self._L.pte_serialize_payloads(self._h, first, count, blob, len(blob), _ptr(off, C.c_uint32), n_phase, n_ns,
                                                None, 0, None, C.byref(need)))
        out = np.empty(max(int(need.value), 1), np.uint8)
        offsets = np.zeros(count + 1, np.uint64)
        self._ck(self._L.pte_serialize_payloads(self._h, first, count, blob, len(blob), _ptr(off, C.c_uint32), n_phase, n_ns,
                                                out.ctypes.data_as(C.c_void_p), out.shape[0], _ptr(offsets, C.c_uint64),
                                                C.byref(need)))
        return out[: int(need.value)], offsets

    def fetch_bitmap(self, into: Optional[np.ndarray] = None) -> np.ndarray:
        """Emit bitmap of the last run: bit i set iff event i of the resident log is notified (pte_fetch_bitmap)."""
        n = int(self.last_stats["n_in"]) if self.last_stats else 0
        words = (n + 31) // 32
        out = into if into is not None else np.zeros(words, np.uint32)
        got = C.c_uint64()
        self._ck(self._L.pte_fetch_bitmap(self._h, _ptr(out, C.c_uint32), int(out.shape[0]), C.byref(got)))
        return out[:words]

    def replay_pipelined(self, cols: Dict[str, np.ndarray], *, dedupe: bool = True, chunk_events: int = 0,
                         with_namespaces: bool = True, with_rv: bool = True, bitmap_out: Optional[np.ndarray] = None,
                         meta: Optional[np.ndarray] = None):
        """ingest + run + fetch_bitmap as one pipelined call (pte_replay_pipelined): host -> device copies of later
        chunks overlap the replay of earlier ones.  Columns in pinned memory (PinnedBuffer) copy at full PCIe speed.
        ``meta``: one byte per event, ``type << 6 | phase`` (pack_meta), sent instead of the phase and type columns.
        Returns (emit bitmap words, stats)."""
        n = int(cols["rv"].shape[0])
        c = {name: np.ascontiguousarray(cols[name], dtype=dt) for name, dt in COLUMNS}
        if meta is not None:
            meta = np.ascontiguousarray(meta, dtype=np.uint8)
            if meta.shape[0] != n:
                raise ValueError("meta has %d rows, the log %d" % (meta.shape[0], n))
        words = (n + 31) // 32
        out = bitmap_out if bitmap_out is not None else np.zeros(max(words, 1), np.uint32)
        st = PteStats()
        self._ck(self._L.pte_replay_pipelined(
            self._h, _ptr(c["uid_hi"], C.c_uint64), _ptr(c["uid_lo"], C.c_uint64),
            _ptr(c["rv"], C.c_uint64) if with_rv else None,      # with_rv=False: the caller vouches for arrival order
            _ptr(c["ns_id"], C.c_uint32) if with_namespaces else None,
            _ptr(c["phase"] if meta is None else meta, C.c_uint8), _ptr(c["type"], C.c_uint8) if meta is None else None,
            n, int(chunk_events), int(bool(dedupe)), _ptr(out, C.c_uint32), C.byref(st)))
        self.last_stats = st.as_dict()
        return out[:words], self.last_stats

    def stream_batch(self, cols: Dict[str, np.ndarray], *, dedupe: bool = True, carry: bool = True):
        """One micro-batch of a live stream in one call (pte_stream_batch): -> (EmittedColumns with idx and type, stats)."""
        n = int(cols["rv"].shape[0])
        c = {name: np.ascontiguousarray(cols[name], dtype=dt) for name, dt in COLUMNS}
        idx, ty = np.empty(n, np.uint32), np.empty(n, np.uint8)
        st = PteStats()
        flags = (1 if dedupe else 0) | (2 if carry and dedupe else 0)
        # the live loop calls this once per batch: plain addresses instead of eight typed ctypes pointers (3.5 us each)
        f = self._stream_batch_fn
        if f is None:
            f = self._stream_batch_fn = self._L["pte_stream_batch"]
            f.restype = C.c_int
            f.argtypes = [C.c_void_p] * 7 + [C.c_uint64, C.c_int, C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(PteStats)]
        self._ck(f(self._h, c["uid_hi"].ctypes.data, c["uid_lo"].ctypes.data, c["rv"].ctypes.data, c["ns_id"].ctypes.data,
                   c["phase"].ctypes.data, c["type"].ctypes.data, n, flags, idx.ctypes.data, ty.ctypes.data, n, C.byref(st)))
        self.last_stats = st.as_dict()
        k = int(st.n_emitted)
        z = np.empty(0, np.uint64)
        return EmittedColumns(idx=idx[:k], rv=z, key=z, ns_id=np.empty(0, np.uint32), phase=np.empty(0, np.uint8), type=ty[:k]), self.last_stats

    def grow(self, max_events: int):
        """Raise max_events in place; dictionary and carried per-pod state survive (pte_grow)."""
        self._ck(self._L.pte_grow(self._h, int(max_events)))
        self.max_events = max(self.max_events, int(max_events))

    def digest(self) -> int:
        d = C.c_uint64()
        self._ck(self._L.pte_digest(self._h, C.byref(d)))
        return int(d.value)

    def debug_sort(self, records: np.ndarray, bit_lo: int, bit_hi: int) -> np.ndarray:
        r = np.ascontiguousarray(records, dtype=np.uint64).copy()
        self._ck(self._L.pte_debug_sort_u64(self._h, _ptr(r, C.c_uint64), int(r.shape[0]), bit_lo, bit_hi))
        return r

    # -- multi-GPU
    @staticmethod
    def comm_unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        _abi.check(_abi.lib().pte_comm_unique_id(buf))
        return buf.raw

    def comm_init(self, unique_id: bytes, rank: int, n_ranks: int):
        assert len(unique_id) == 128
        self.n_ranks = n_ranks
        self._ck(self._L.pte_comm_init(self._h, unique_id, rank, n_ranks))

    def gather_counts(self) -> np.ndarray:
        n = getattr(self, "n_ranks", 1)
        out = np.zeros(4 * n, np.uint64)
        self._ck(self._L.pte_gather_counts(self._h, _ptr(out, C.c_uint64)))
        return out.reshape(n, 4)
