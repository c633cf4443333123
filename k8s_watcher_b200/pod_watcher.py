"""Drop-in for the reference's ``watcher.pod_watcher.PodWatcher`` whose event path runs in bulk on a B200.

Same class surface as the reference (watcher/pod_watcher.py:10-277): ``PodWatcher(environment)``
with ``.environment .config .v1 .watch .logger``, ``setup_k8s_client()``,
``should_process_event()``, ``handle_pod_event()``, ``_extract_pod_data()``, ``start_watching()``,
and the same error conventions (set-up failure -> False + error log, stream failure -> log and
re-raise, ``watch.stop()`` always).  What changes is where the work happens:

* ``handle_pod_event`` does not decide anything itself; it queues the event.
* ``flush()`` / ``replay(log)`` hand the whole batch to the CUDA engine (libpte.so), which applies
  the reference's two filters, and -- with ``dedupe=True`` -- the last-known-phase layer the
  reference's README promises but pod_watcher.py never implemented.
  ``dedupe=False`` reproduces the reference literally: every event that reaches the POST site.
* emitted records come back as a :class:`Batch`; the sink takes it in one call.

DEVIATION FROM THE REFERENCE, on purpose: the default is ``dedupe=True`` (transition notifications only).
The reference as shipped would notify on EVERY event that passes its filters; construct the watcher
with ``dedupe=False`` to get exactly that.  The live loop flushes every ``flush_every`` events
(default 4096) or ``flush_interval`` seconds (default 1.0), whichever comes first, and on every exit path.
Events are ordered by ARRIVAL, like the reference's ``for event in stream`` loop: the live loop numbers
them itself and never looks at ``metadata.resource_version`` (the reference does not read it either).

There is no CPU implementation of the path in this package: without libpte.so and a CUDA
device ``replay``/``flush`` raise.
"""
from __future__ import annotations

import logging
import os
import time
from datetime import datetime
from typing import Any, Dict, Iterator, List, Optional

import yaml

try:
    from kubernetes import client, config, watch
    from kubernetes.config import ConfigException
except ImportError:                      # no kubernetes package: replay() still works, the live loop reports the set-up failure
    client = config = watch = None

    class ConfigException(Exception):
        pass

from . import codes
from .clusterapi_client import ClusterApiClient
from .engine import EmittedColumns, PhaseTransitionEngine, PteError
from .replay_log import ReplayLog


class _RawJsonLog:
    """What Batch needs of a log, for a chunk of raw watch lines: the pod object of line i, made on demand."""

    def __init__(self, text: bytes, n: int, deserialize):
        self._lines = text.split(b"\n")
        self._n, self._deserialize = n, deserialize
        self._pods = {}              # (not None: bodies are built from the objects, not on the device)
        self.deserialized = 0

    def __len__(self):
        return self._n

    def pod_at(self, i: int):
        pod = self._pods.get(i)
        if pod is None:
            pod = self._pods[i] = self._deserialize(self._lines[i])
            self.deserialized += 1
        return pod


class Batch:
    """Emitted notifications of one replay, in replay order."""

    def __init__(self, watcher: "PodWatcher", log: ReplayLog, emitted: Optional[EmittedColumns], stats: Dict[str, float],
                 dedupe: bool, bitmap=None):
        self.watcher, self.log, self.stats, self.dedupe = watcher, log, stats, dedupe
        self._emitted, self.bitmap = emitted, bitmap      # large replays come back as one bit per event (pte_fetch_bitmap)

    @property
    def emitted(self) -> EmittedColumns:
        """Arrival index and event type of every notification (expanded from the emit bitmap on first use)."""
        if self._emitted is None:
            import numpy as np
            n = len(self.log)
            idx = np.flatnonzero(np.unpackbits(self.bitmap.view(np.uint8), bitorder="little")[:n]).astype(np.uint32)
            z = np.empty(0, np.uint64)
            self._emitted = EmittedColumns(idx=idx, rv=z, key=z, ns_id=np.empty(0, np.uint32), phase=np.empty(0, np.uint8),
                                           type=self.log.columns["type"][idx])
        return self._emitted

    def __len__(self):
        return int(self.stats["n_emitted"]) if self._emitted is None else len(self._emitted)

    @property
    def indices(self):
        """Arrival indices (into the log) of the emitted events."""
        return self.emitted.idx

    def wire_bodies(self, event_timestamp: Optional[str] = None) -> List[bytes]:
        """The request bodies of all notifications, as ``requests`` would send them.  For a column log
        (rows only, no pod objects) they are serialised on the device in one call; otherwise per payload."""
        ts = event_timestamp if event_timestamp is not None else datetime.now().isoformat()
        eng = self.watcher._engine
        if (self.log._pods is None and eng is not None and self.bitmap is None and self is self.watcher.batches[-1]
                and len(self)):
            frags, n_phase, n_ns = payload_fragments(self.watcher, self.log, ts)
            buf, off = eng.serialize_payloads(frags, n_phase, n_ns)
            raw = buf.tobytes()
            return [raw[int(off[i]):int(off[i + 1])] for i in range(len(self))]
        import json
        return [json.dumps(p, allow_nan=False).encode("utf-8") for p in self.payloads(event_timestamp=ts)]

    def payloads(self, event_timestamp: Optional[str] = None) -> Iterator[Dict[str, Any]]:
        """Payload dicts exactly as the reference builds them at pod_watcher.py:232-233,
        one per emitted record.  ``event_timestamp`` pins the wall-clock field (:199)."""
        for i, t in zip(self.emitted.idx.tolist(), self.emitted.type.tolist()):
            d = self.watcher._extract_pod_data(self.log.pod_at(i), event_timestamp=event_timestamp)
            d["event_type"] = codes.TYPE_NAMES[t]
            yield d


def payload_fragments(watcher: "PodWatcher", log: ReplayLog, event_timestamp: str):
    """JSON text around the per-record fields of a column log's request body, for pte_serialize_payloads.

    Derived from json.dumps of a sample payload with placeholders, so key order, separators and escaping
    are whatever the reference's ``session.post(json=...)`` produces (clusterapi_client.py:36).
    Returns (fragments, n_phase, n_ns)."""
    import json
    marks = {k: "\u0000%s\u0000" % k for k in ("NAME", "NS", "UID", "PHASE", "TYPE")}

    class _P:  # a pod whose per-record fields are placeholders
        pass
    pod = log.pod_at(0) if len(log) else None
    md = type("M", (), dict(name="pod-" + marks["NAME"], namespace=marks["NS"], uid=marks["UID"], labels=None,
                            annotations=None, creation_timestamp=None))()
    st = type("S", (), dict(phase=marks["PHASE"], conditions=None, container_statuses=None))()
    sample = watcher._extract_pod_data(type("P", (), dict(metadata=md, status=st, spec=None))(),
                                       event_timestamp=event_timestamp)
    sample["event_type"] = marks["TYPE"]
    text = json.dumps(sample, allow_nan=False)
    # split at the (escaped) placeholders; quotes around NS / PHASE / TYPE belong to the fragment tables
    def esc(m):
        return json.dumps(m)[1:-1]
    parts, rest = [], text
    for key, with_quotes in (("NAME", False), ("NS", True), ("UID", False), ("PHASE", True), ("TYPE", True)):
        token = ('"%s"' % esc(marks[key])) if with_quotes else esc(marks[key])
        head, rest = rest.split(token, 1)
        parts.append(head)
    parts.append(rest)
    segs = [p.encode("utf-8") for p in parts]
    types = [json.dumps(t).encode() for t in codes.TYPE_NAMES]
    phases = []
    for code in range(max(len(log.phases), codes.PH_NOSTATUS + 1)):
        if code == codes.PH_NOSTATUS:
            phases.append(b'"Unknown"')                       # pod.status is None (pod_watcher.py:167)
        elif code == codes.PH_NULL or code >= len(log.phases) or log.phases[code] is None:
            phases.append(b"null")
        else:
            phases.append(json.dumps(log.phases[code]).encode())
    nss = [json.dumps(n).encode() for n in log.namespaces] or [b'""']
    return segs + types + phases + nss, len(phases), len(nss)


class PodWatcher:
    def __init__(self, environment: str = "development", *, dedupe: bool = True,
                 sink: Optional[ClusterApiClient] = None, engine: Optional[PhaseTransitionEngine] = None,
                 device: int = 0, config_dir: str = "config", flush_every: Optional[int] = 4096,
                 flush_interval: Optional[float] = 1.0, max_pods: int = 1 << 20, keep_batches: Optional[int] = 4):
        self.environment = environment
        self.config_dir = config_dir           # CWD-relative by default, like upstream (:22,:25)
        self.config = self._load_environment_config()
        self.v1 = None
        self.watch = watch.Watch() if watch is not None else None
        self._setup_logging()
        self.dedupe = dedupe
        self.clusterapi_client = sink          # the reference leaves this commented out (:14)
        self._engine = engine
        self._device = device
        self._pending: List[Dict[str, Any]] = []
        # every replay's result, oldest first; a watcher that runs for days sets keep_batches so that only the
        # last few (each holds its log) stay alive
        self.batches: List[Batch] = []
        self.keep_batches = keep_batches
        # live loop: replay every `flush_every` queued events as one batch of a stream; the engine carries the
        # uid dictionary and each pod's last known phase from batch to batch (None: one replay at stream end)
        self.pipeline_min_events = 1 << 22          # replay() pipelines logs at least this long (pte_replay_pipelined)
        self.pipeline_chunk_events = 1 << 26
        self.flush_every = flush_every
        self.flush_interval = flush_interval
        self._last_flush = time.monotonic()
        self._seq = 0                                # arrival number of the next live event (the stream's order key)
        self.max_pods = max_pods
        self._stream_namespaces: List[str] = []      # namespace dictionary shared by all batches of a stream
        self._stream_uids = codes.UidInterner()      # uid string -> 128-bit key, shared by all batches of a stream

    # ---- configuration (pod_watcher.py:19-75) ---------------------------------
    def _load_environment_config(self) -> Dict[str, Any]:
        base = self._load_config_file("%s/base.yaml" % self.config_dir)
        env = self._load_config_file("%s/%s.yaml" % (self.config_dir, self.environment))
        return self._substitute_env_vars(self._merge_configs(base, env))

    def _load_config_file(self, config_file: str) -> Dict[str, Any]:
        def say(msg, level):
            if hasattr(self, "logger"):
                getattr(self.logger, level)(msg)
            else:
                print(msg)
        try:
            with open(config_file, "r") as fh:
                return yaml.safe_load(fh) or {}
        except FileNotFoundError:
            say("Config file %s not found" % config_file, "warning")
            return {}
        except Exception as e:
            say("Error loading config %s: %s" % (config_file, e), "error")
            return {}

    def _merge_configs(self, base: Dict[str, Any], override: Dict[str, Any]) -> Dict[str, Any]:
        merged = dict(base)
        for key, value in override.items():
            if isinstance(merged.get(key), dict) and isinstance(value, dict):
                merged[key] = self._merge_configs(merged[key], value)
            else:
                merged[key] = value
        return merged

    def _substitute_env_vars(self, cfg: Dict[str, Any]) -> Dict[str, Any]:
        def walk(node):
            if isinstance(node, dict):
                return {k: walk(v) for k, v in node.items()}
            if isinstance(node, list):
                return [walk(v) for v in node]
            if isinstance(node, str) and node.startswith("${") and node.endswith("}"):
                name, default = node[2:-1], ""
                if ":-" in name:
                    name, default = name.split(":-", 1)
                return os.getenv(name, default)
            return node
        return walk(cfg)

    def _setup_logging(self):
        level = getattr(logging, self.config.get("watcher", {}).get("log_level", "INFO"))
        if self.environment == "production":          # JSON-shaped lines in production (:82-84)
            fmt = ('{"timestamp":"%(asctime)s","level":"%(levelname)s","logger":"%(name)s",'
                   '"message":"%(message)s","environment":"' + self.environment + '"}')
        else:
            fmt = "[%s] %%(asctime)s - %%(name)s - %%(levelname)s - %%(message)s" % self.environment.upper()
        logging.basicConfig(level=level, format=fmt)
        self.logger = logging.getLogger(__name__)
        self.logger.info("Starting k8s-watcher in %s environment", self.environment)

    # ---- kubernetes client (pod_watcher.py:110-157) -----------------------------
    def setup_k8s_client(self) -> bool:
        if client is None:
            self.logger.error("Error setting up k8s client: the kubernetes package is not installed")
            return False
        try:
            k8s = self.config.get("kubernetes", {})
            if k8s.get("use_incluster_config", False):
                self.logger.info("Using in-cluster configuration")
                config.load_incluster_config()
            elif "config_file" in k8s:
                path = k8s["config_file"]
                self.logger.info("Loading kubeconfig from: %s", path)
                if not os.path.exists(path):
                    self.logger.error("Kubeconfig file not found: %s", path)
                    return False
                config.load_kube_config(config_file=path)
            else:
                self.logger.info("Using default kubeconfig")
                config.load_kube_config()
            self.v1 = client.CoreV1Api()
            self.logger.info("Successfully connected to Kubernetes API version: %s", self.v1.get_api_version())
            if self.environment != "local":
                sample = [ns.metadata.name for ns in self.v1.list_namespace().items[:5]]
                self.logger.info("Sample namespaces: %s", sample)
            return True
        except ConfigException as e:
            self.logger.error("Kubernetes config error: %s", e)
            return False
        except Exception as e:
            self.logger.error("Error setting up k8s client: %s", e)
            return False

    # ---- per-event surface, kept for callers and for payload materialisation -----
    def _extract_pod_data(self, pod, event_timestamp: Optional[str] = None) -> Dict[str, Any]:
        """Payload for one pod; same keys, order and defaults as pod_watcher.py:159-202."""
        md, st, sp = pod.metadata, pod.status, pod.spec
        conditions, container_statuses, containers = [], [], []
        if st:
            for c in (st.conditions or []):
                conditions.append({"type": c.type, "status": c.status, "reason": c.reason, "message": c.message})
            for cs in (st.container_statuses or []):
                container_statuses.append({"name": cs.name, "ready": cs.ready, "restart_count": cs.restart_count,
                                           "state": str(cs.state) if cs.state else None})
        if sp:
            for c in (sp.containers or []):
                containers.append({"name": c.name, "image": c.image})
        created = md.creation_timestamp
        return {
            "name": md.name,
            "namespace": md.namespace,
            "uid": md.uid,
            "environment": self.environment,
            "status": {"phase": st.phase if st else "Unknown", "conditions": conditions,
                       "container_statuses": container_statuses},
            "spec": {"node_name": sp.node_name if sp else None, "containers": containers},
            "metadata": {"labels": md.labels or {}, "annotations": md.annotations or {},
                         "creation_timestamp": created.isoformat() if created else None},
            "event_timestamp": event_timestamp if event_timestamp is not None else datetime.now().isoformat(),
        }

    def _critical_only(self) -> bool:
        return self.environment == "production" and bool(
            self.config.get("watcher", {}).get("alerts", {}).get("critical_events_only", False))

    def should_process_event(self, event_type: str, pod) -> bool:
        """Host form of the first device predicate (pod_watcher.py:204-212)."""
        if self._critical_only() and event_type not in ["DELETED"] and pod.status \
                and pod.status.phase not in ["Failed", "Succeeded"]:
            return False
        return True

    def handle_pod_event(self, event_type: str, pod) -> None:
        """Queue one event for the next bulk replay (the reference decides per event, :214-241)."""
        md = getattr(pod, "metadata", None)
        self.logger.info("Pod event detected: %s - %s/%s", event_type, getattr(md, "namespace", None), getattr(md, "name", None))
        self._pending.append({"type": event_type, "object": pod})
        due = self.flush_interval is not None and time.monotonic() - self._last_flush >= self.flush_interval
        if (self.flush_every and len(self._pending) >= self.flush_every) or due:
            self.flush()

    # ---- bulk path -----------------------------------------------------------------
    def _engine_for(self, n_events: int, n_pods_hint: int) -> PhaseTransitionEngine:
        eng = self._engine
        if eng is not None and eng.max_events < n_events and (n_pods_hint or n_events) <= eng.max_pods:
            eng.grow(n_events)           # same dictionary, larger per-event buffers: no destroy-and-recreate
        elif eng is None or eng.max_events < n_events or n_pods_hint > eng.max_pods:
            if eng is not None:
                eng.close()
            eng = self._engine = PhaseTransitionEngine(max(n_events, 1024), max(n_pods_hint, 0), device=self._device)
        return eng

    def replay(self, log: ReplayLog, *, dedupe: Optional[bool] = None, max_pods: int = 0, carry: bool = False) -> Batch:
        """Run a whole log through the engine and return what a watcher would have notified.
        ``carry=True`` treats the log as the next batch of a stream (state kept on the device)."""
        dedupe = self.dedupe if dedupe is None else dedupe
        if carry and (self._engine is None or self._engine.max_events < len(log)):
            # the dictionary is sized by max_pods, independently of the batch size; a longer batch than ever before
            # moves the stream to a larger engine (pte_grow keeps dictionary and per-pod state)
            want = max(self.flush_every or 0, len(log), 1024)
            if self._engine is None:
                self._engine = PhaseTransitionEngine(want, self.max_pods, device=self._device)
            else:
                self._engine.grow(want)
        eng = self._engine if carry else self._engine_for(len(log), max_pods)
        names = self.config.get("watcher", {}).get("namespaces", []) or []
        eng.set_filters(log.namespace_ids(names), ns_filter_on=bool(names), critical_only=self._critical_only())
        if not carry and len(log) >= self.pipeline_min_events:
            # a long log: copies, replay and the way back overlap chunk by chunk, and the result is one bit per event
            # (the host holds the log, so that is all the payload hand-off needs); the namespace column stays at
            # home when no allow-list could look at it
            try:
                bitmap, stats = eng.replay_pipelined(log.columns, dedupe=dedupe, with_namespaces=bool(names),
                                                     with_rv=not log.arrival_ordered,
                                                     chunk_events=self.pipeline_chunk_events,
                                                     bitmap_out=log.bitmap_buffer(), meta=log.meta)
                batch = Batch(self, log, None, stats, dedupe, bitmap=bitmap)
                self._remember(batch)
                return batch
            except PteError as e:
                if e.code != -7:                  # PTE_ERR_UNORDERED: the plain path pre-sorts such a log
                    raise
        if carry:
            # a micro-batch of the live loop: copy in, replay, notifications out in one call
            emitted, stats = eng.stream_batch(log.columns, dedupe=dedupe, carry=True)
            batch = Batch(self, log, emitted, stats, dedupe)
            self._remember(batch)
            return batch
        eng.ingest(log.columns)          # with the namespace column: Batch.wire_bodies may serialise payloads on the device
        stats = eng.run(dedupe=dedupe, carry=carry)
        # the host holds the log, so the arrival index (+ event type) is all the payload hand-off needs
        batch = Batch(self, log, eng.fetch(columns=("idx", "type")), stats, dedupe)
        self._remember(batch)
        return batch

    def _remember(self, batch: Batch) -> None:
        self.batches.append(batch)
        if self.keep_batches is not None and len(self.batches) > max(1, self.keep_batches):
            del self.batches[:len(self.batches) - max(1, self.keep_batches)]

    def replay_watch_json(self, text: bytes, *, dedupe: Optional[bool] = None, max_events: int = 0, max_pods: int = 0) -> Batch:
        """Replay a raw watch log (newline-delimited JSON watch events, what ``for event in stream`` would have
        yielded, :264-268): parsed, interned and replayed on the device."""
        dedupe = self.dedupe if dedupe is None else dedupe
        n_lines = text.count(b"\n") + 1
        eng = self._engine_for(max(max_events, n_lines), max_pods)
        log = ReplayLog.from_watch_json(eng, text)
        names = self.config.get("watcher", {}).get("namespaces", []) or []
        eng.set_filters(log.namespace_ids(names), ns_filter_on=bool(names), critical_only=self._critical_only())
        stats = eng.run(dedupe=dedupe)
        batch = Batch(self, log, eng.fetch(columns=("idx", "type")), stats, dedupe)
        self._remember(batch)
        return batch

    def watch_json(self, chunks, deserialize, *, dedupe: Optional[bool] = None) -> List[Batch]:
        """The live loop over the RAW watch stream (what the API server sends for ``?watch=true``, before any client has
        turned it into objects): ``chunks`` yields bytes, each a run of whole lines, one JSON watch event per line, in
        arrival order.  Every chunk is parsed, filtered and compared with the stream's per-pod state on the device
        (pte_json_stream + PTE_RUN_CARRY); ``deserialize(line) -> pod object`` is called ONLY for the lines that lead to
        a notification -- with the kubernetes client that is the expensive step of ``for event in w.stream(...)``
        (pod_watcher.py:264-269), and under the production filter most events never need it.  Payloads and sink calls are
        those of the object loop.  Returns the batches."""
        dedupe = self.dedupe if dedupe is None else dedupe
        names = self.config.get("watcher", {}).get("namespaces", []) or []
        out: List[Batch] = []
        try:
            return self._watch_json(chunks, deserialize, dedupe, names, out)
        finally:
            if self._engine is not None:
                self._engine.json_stream(False)          # whatever ended the loop, the next log is its own log again

    def _watch_json(self, chunks, deserialize, dedupe, names, out) -> List[Batch]:
        started = False
        for chunk in chunks:
            if not chunk.strip():
                continue
            n_lines = chunk.count(b"\n") + (0 if chunk.endswith(b"\n") else 1)
            if self._engine is None:
                self._engine = PhaseTransitionEngine(max(self.flush_every or 0, n_lines, 1024), self.max_pods, device=self._device)
            elif self._engine.max_events < n_lines:
                self._engine.grow(n_lines)
            eng = self._engine
            if not started:
                eng.json_stream(True)
                started = True
            n = eng.ingest_json(chunk)
            known = eng.json_strings(0)                    # the stream's namespaces so far, index = ns_id
            eng.set_filters([i for i, nm in enumerate(known) if nm in names], ns_filter_on=bool(names),
                            critical_only=self._critical_only())
            stats = eng.run(dedupe=dedupe, carry=True)
            batch = Batch(self, _RawJsonLog(chunk, n, deserialize), eng.fetch(columns=("idx", "type")), stats, dedupe)
            self._remember(batch)
            out.append(batch)
            if self.clusterapi_client is not None:
                sent = self.clusterapi_client.update_pod_status_batch(batch)
                if sent != len(batch):
                    self.logger.error("Failed to notify clusterapi about %d of %d events", len(batch) - sent, len(batch))
        return out

    def _stream_mode(self) -> bool:
        return bool(self.flush_every) or self.flush_interval is not None

    def flush(self) -> Optional[Batch]:
        """Replay everything queued by handle_pod_event; hand the result to the sink if there is one."""
        if not self._pending:
            return None
        # the stream's order key is the arrival number, as in the reference's loop (metadata.resource_version is
        # neither read upstream nor ordered across the initial list of a list+watch)
        log = ReplayLog.from_events(self._pending, namespaces=self._stream_namespaces, seq_start=self._seq + 1,
                                    uids=self._stream_uids)
        self._seq += len(self._pending)
        self._pending = []
        self._last_flush = time.monotonic()
        batch = self.replay(log, carry=self._stream_mode())
        if self.clusterapi_client is not None:
            sent = self.clusterapi_client.update_pod_status_batch(batch)
            if sent != len(batch):
                self.logger.error("Failed to notify clusterapi about %d of %d events", len(batch) - sent, len(batch))
        return batch

    def start_watching(self):
        """Live loop (pod_watcher.py:243-277): same set-up, same error handling; events are queued
        and replayed in bulk when the stream ends or the watcher is interrupted."""
        if not self.setup_k8s_client():
            self.logger.error("Failed to setup Kubernetes client")
            return
        # the health gate the reference has commented out (:249-253): warn and carry on
        if self.clusterapi_client is not None:
            if not self.clusterapi_client.health_check():
                self.logger.warning("ClusterAPI health check failed, but continuing...")
            else:
                self.logger.info("ClusterAPI health check passed")

        self.logger.info("Starting Pod watcher in %s environment...", self.environment)
        targets = self.config.get("watcher", {}).get("namespaces", [])
        if targets:
            self.logger.info("Monitoring namespaces: %s", targets)
        else:
            self.logger.info("Monitoring all namespaces")
        try:
            for event in self.watch.stream(self.v1.list_pod_for_all_namespaces):
                self.handle_pod_event(event["type"], event["object"])
        except KeyboardInterrupt:
            self.logger.info("Stopping Pod watcher...")
        except Exception as e:
            self.logger.error("Error in Pod watcher: %s", e)
            raise
        finally:
            try:
                self.flush()             # whatever is still queued goes out on every exit path
            finally:
                self.watch.stop()
