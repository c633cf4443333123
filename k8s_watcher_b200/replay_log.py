"""Replay logs: pod-watch events as integer columns plus the dictionaries to get back.

A log row is what the reference reads from one watch event (watcher/pod_watcher.py:162-167,
:216-217, :267-268) plus ``metadata.resource_version`` as the order key.
"""
from __future__ import annotations

import datetime as _dt
from typing import Any, Dict, Iterable, List, Optional

import numpy as np

from . import codes
from .engine import COLUMNS, generate_host


class _Attr:
    def __init__(self, **kw):
        self.__dict__.update(kw)


class ReplayLog:
    def __init__(self, columns: Dict[str, np.ndarray], namespaces: List[str], phases: Optional[List[Any]] = None,
                 pods: Optional[List[Any]] = None):
        self.columns = {name: np.ascontiguousarray(columns[name], dtype=dt) for name, dt in COLUMNS}
        self.namespaces = list(namespaces)
        # code -> phase string; codes 0..6 are fixed by include/pte.h
        self.phases = list(phases) if phases else list(codes.PHASE_NAMES) + [None, None]
        self._pods = pods            # per-row pod objects when the log came from a live stream
        # True: the rows are in the order the watcher received them and that IS the replay order (the reference's loop
        # knows no other; it never reads resourceVersion).  replay() then leaves the rv column on the host.
        self.arrival_ordered = False
        self.meta = None             # pack_meta(): type << 6 | phase per event
        self.uids: Optional[codes.UidInterner] = None      # set by from_events: maps keys of non-UUID uids back

    def __len__(self):
        return int(self.columns["rv"].shape[0])

    # ---- page-locked staging: what a host that replays big logs wants (full PCIe speed, reusable result buffer) ----
    @classmethod
    def pinned(cls, n_events: int, namespaces: List[str]) -> "ReplayLog":
        """An empty log of n_events rows whose columns (and emit-bitmap buffer) live in pinned host memory
        (pte_host_alloc); fill ``log.columns[...]`` in place."""
        from .engine import PinnedBuffer
        bufs = {name: PinnedBuffer(n_events, dt) for name, dt in COLUMNS}
        log = cls({name: b.array for name, b in bufs.items()}, namespaces)
        log._pins = bufs
        log._bitmap_pin = PinnedBuffer((n_events + 31) // 32 + 1, np.uint32)
        return log

    def pack_meta(self):
        """Keep a packed (type, phase) byte per event beside the columns (pinned if the log is): replay() then sends that one
        byte instead of two.  To be called again after phase or type columns change."""
        from .engine import PinnedBuffer, pack_meta
        n = len(self)
        if getattr(self, "_pins", None):
            pin = self._pins.get("meta")
            if pin is None:
                pin = self._pins["meta"] = PinnedBuffer(n, np.uint8)
            self.meta = pack_meta(self.columns["type"], self.columns["phase"], out=pin.array)
        else:
            self.meta = pack_meta(self.columns["type"], self.columns["phase"])
        return self.meta

    def bitmap_buffer(self):
        """Reusable destination of the emit bitmap (pinned if the log is), or None."""
        pin = getattr(self, "_bitmap_pin", None)
        return pin.array if pin is not None else None

    def free(self):
        for b in list(getattr(self, "_pins", {}).values()) + [getattr(self, "_bitmap_pin", None)]:
            if b is not None:
                b.free()
        self._pins, self._bitmap_pin = {}, None

    # ---- construction -------------------------------------------------------
    @classmethod
    def from_events(cls, events: Iterable[Dict[str, Any]], namespaces: Optional[List[str]] = None,
                    seq_start: Optional[int] = None, uids: Optional[codes.UidInterner] = None) -> "ReplayLog":
        """Intern a stream of ``{'type': str, 'object': V1Pod}`` dicts (pod_watcher.py:267-268).
        ``namespaces`` is an optional, growing id -> name list shared between the batches of one stream.
        ``seq_start``: order the events by arrival (``rv = seq_start + i``) instead of by their
        ``metadata.resource_version`` -- what a live watch loop wants, see PodWatcher.flush.
        ``uids``: the stream's UidInterner (exact keys for uid strings of any form, stable across batches); a fresh one
        per log otherwise."""
        events = list(events)
        uids = uids if uids is not None else codes.UidInterner()
        n = len(events)
        cols = {name: np.zeros(n, dt) for name, dt in COLUMNS}
        shared = namespaces if namespaces is not None else []
        ns_ids: Dict[str, int] = {name: i for i, name in enumerate(shared)}
        phase_ids: Dict[Any, int] = {p: i for i, p in enumerate(codes.PHASE_NAMES)}
        phases: List[Any] = list(codes.PHASE_NAMES) + [None, None]
        pods = []
        for i, ev in enumerate(events):
            t = ev["type"]
            pod = ev["object"]
            if t not in codes.TYPE_CODE:
                raise ValueError("event %d: type %r is not ADDED/MODIFIED/DELETED" % (i, t))
            md = pod.metadata
            hi, lo = uids.key(md.uid)
            cols["uid_hi"][i], cols["uid_lo"][i] = hi, lo
            rvs = getattr(md, "resource_version", None)
            if seq_start is not None:
                cols["rv"][i] = seq_start + i
            else:
                cols["rv"][i] = int(rvs) if rvs not in (None, "") else i + 1
            cols["ns_id"][i] = ns_ids.setdefault(md.namespace, len(ns_ids))
            st = pod.status
            if not st:
                code = codes.PH_NOSTATUS
            elif st.phase is None:
                code = codes.PH_NULL
            else:
                code = phase_ids.get(st.phase)
                if code is None:
                    code = len(phases)
                    if code > codes.PH_MAX:
                        raise ValueError("more than %d distinct phase strings" % (codes.PH_MAX - 6))
                    phase_ids[st.phase] = code
                    phases.append(st.phase)
            cols["phase"][i] = code
            cols["type"][i] = codes.TYPE_CODE[t]
            pods.append(pod)
        names = [None] * len(ns_ids)
        for k, v in ns_ids.items():
            names[v] = k
        shared[:] = names
        log = cls(cols, names, phases, pods)
        log.uids = uids
        log.arrival_ordered = seq_start is not None
        return log

    @classmethod
    def from_watch_json(cls, engine, text: bytes) -> "ReplayLog":
        """Raw watch log (newline-delimited JSON watch events) -> replay log, parsed and interned on the device
        (pte_ingest_json).  The log is left resident in ``engine``, ready for ``engine.run()``."""
        engine.ingest_json(text)
        phases = list(codes.PHASE_NAMES) + [None, None] + engine.json_strings(1)
        log = cls(engine.read_log(), engine.json_strings(0), phases)
        odd = engine.json_strings(2)
        if odd:                       # uid strings that are not canonical UUIDs: the device's keys for them, and the way back
            log.uids = codes.UidInterner()
            for name in odd:
                log.uids._ids[name] = len(log.uids._strings) + 1
                log.uids._strings.append(name)
        return log

    @classmethod
    def synthetic(cls, seed: int, n_events: int, n_pods: int, **kw) -> "ReplayLog":
        """The synthetic log of include/pte.h (host generator of libpte.so)."""
        return cls(generate_host(seed, n_events, n_pods, **kw), list(codes.DEFAULT_NAMESPACES))

    # ---- the way back ---------------------------------------------------------
    def namespace_ids(self, names: Iterable[str]) -> List[int]:
        """ids of those allow-list names that occur in this log (others can never match)."""
        lut = {n: i for i, n in enumerate(self.namespaces)}
        return [lut[n] for n in names if n in lut]

    def pod_at(self, i: int):
        """Pod object of row i: the original one for stream logs, a minimal one for column logs."""
        if self._pods is not None:
            return self._pods[i]
        c = self.columns
        code = int(c["phase"][i])
        uid = self.uids.string(c["uid_hi"][i], c["uid_lo"][i]) if self.uids else codes.uid_to_str(c["uid_hi"][i], c["uid_lo"][i])
        md = _Attr(name="pod-" + uid[:13], namespace=self.namespaces[int(c["ns_id"][i])], uid=uid, labels=None,
                   annotations=None, creation_timestamp=None, resource_version=str(int(c["rv"][i])))
        status = None if code == codes.PH_NOSTATUS else _Attr(phase=self.phases[code], conditions=None,
                                                              container_statuses=None)
        return _Attr(metadata=md, status=status, spec=None)

    def type_at(self, i: int) -> str:
        return codes.TYPE_NAMES[int(self.columns["type"][i])]

    def events(self):
        for i in range(len(self)):
            yield {"type": self.type_at(i), "object": self.pod_at(i)}

    def shard(self, n_shards: int) -> List["ReplayLog"]:
        """Split by uid hash (pte_shard_of), keeping arrival order inside every shard."""
        from .engine import shard_of
        hi = self.columns["uid_hi"]
        lo = self.columns["uid_lo"]
        owner = np.fromiter((shard_of(int(h), int(l), n_shards) for h, l in zip(hi, lo)), np.int64, len(self))
        out = []
        for g in range(n_shards):
            sel = np.nonzero(owner == g)[0]
            pods = [self._pods[i] for i in sel] if self._pods is not None else None
            out.append(ReplayLog({k: v[sel] for k, v in self.columns.items()}, self.namespaces, self.phases, pods))
        return out
