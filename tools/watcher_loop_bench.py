"""The drop-in PodWatcher's live loop end to end on this host + GPU: kubernetes-style pod objects in, notifications out
(into an in-memory sink), against the staged reference watcher's own loop on the same objects when oracle/_ref is there.
Where the time goes: turning pod objects into columns (Python, per event) vs the engine call (per batch).

    python tools/watcher_loop_bench.py [events, default 200000] [flush_every, default 4096]
"""
import logging, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import fake_kubernetes as fk
fk.install()
from mock_transport import MockTransport
from k8s_watcher_b200 import ClusterApiClient, PodWatcher, ReplayLog
from oracle import c_oracle

n = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
flush = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
logging.disable(logging.CRITICAL)
os.chdir(ROOT)
cols = c_oracle.generate(11, n, 3000, dist=1, p_del=0.05, p_nostatus=0.02, p_null=0.02)
events = fk.events_from_columns(cols)

fk.set_stream(events)
sink = ClusterApiClient("http://clusterapi.invalid", transport=MockTransport())
w = PodWatcher("staging", sink=sink, flush_every=flush, flush_interval=None, max_pods=1 << 16, keep_batches=2)
t0 = time.perf_counter()
w.start_watching()
t1 = time.perf_counter()
sent = len(sink.transport.bodies())
print("PodWatcher.start_watching: %d events -> %d notifications in %.3f s = %.0f events/s (flush_every=%d)" % (n, sent, t1 - t0, n / (t1 - t0), flush))

# the same loop split up: objects -> columns, engine, payloads + sink
t0 = time.perf_counter()
logs = [ReplayLog.from_events(events[i:i + flush], seq_start=i + 1) for i in range(0, n, flush)]
t1 = time.perf_counter()
print("  ReplayLog.from_events alone (Python, per event): %.3f s = %.0f events/s" % (t1 - t0, n / (t1 - t0)))
from k8s_watcher_b200 import PhaseTransitionEngine
eng = PhaseTransitionEngine(flush, 1 << 16)
t0 = time.perf_counter()
emitted = 0
for lg in logs:
    out, st = eng.stream_batch(lg.columns)
    emitted += len(out.idx)
t1 = time.perf_counter()
print("  engine calls alone (pte_stream_batch per batch): %.3f s = %.0f events/s, %d notifications" % (t1 - t0, n / (t1 - t0), emitted))

# ---- the raw watch stream: PodWatcher.watch_json (objects only for notified lines) against the object loop --------------
import datetime as _dt
import json


def pod_from_line(line: bytes):
    """What a client library does for every event of `for event in w.stream(...)`: JSON text -> pod object
    (here: the fields the payload reads, into the stand-in model classes)."""
    o = json.loads(line)["object"]
    md, st, sp = o.get("metadata") or {}, o.get("status"), o.get("spec")
    ts = md.get("creationTimestamp")
    meta = fk.V1ObjectMeta(name=md.get("name"), namespace=md.get("namespace"), uid=md.get("uid"), labels=md.get("labels"),
                           annotations=md.get("annotations"), resource_version=md.get("resourceVersion"),
                           creation_timestamp=_dt.datetime.fromisoformat(ts[:-1]) if ts else None)
    status = None
    if st:
        status = fk.V1PodStatus(phase=st.get("phase"), container_statuses=None,
                                conditions=[fk.V1PodCondition(type=c.get("type"), status=c.get("status"), reason=None, message=None)
                                            for c in st.get("conditions") or []] or None)
    spec = None if sp is None else fk.V1PodSpec(node_name=sp.get("nodeName"),
                                                containers=[fk.V1Container(name=c.get("name"), image=c.get("image")) for c in sp.get("containers") or []] or None)
    return fk.V1Pod(metadata=meta, status=status, spec=spec)


text = fk.watch_json_lines(events, seed=2)
lines = [l for l in text.split(b"\n") if l]
for envname in ("staging", "production"):
    # the object loop: every line becomes an object, then handle_pod_event queues it and flush() replays the batch
    sink = ClusterApiClient("http://clusterapi.invalid", transport=MockTransport())
    w1 = PodWatcher(envname, sink=sink, flush_every=flush, flush_interval=None, max_pods=1 << 16, keep_batches=2)
    t0 = time.perf_counter()
    for l in lines:
        w1.handle_pod_event(json.loads(l)["type"], pod_from_line(l))
    w1.flush()
    t1 = time.perf_counter()
    a = len(sink.transport.bodies())
    # the raw loop: the GPU reads the text, objects are made for the notified lines only
    sink = ClusterApiClient("http://clusterapi.invalid", transport=MockTransport())
    w2 = PodWatcher(envname, sink=sink, max_pods=1 << 16, keep_batches=2)
    chunks = [b"\n".join(lines[i:i + flush]) for i in range(0, n, flush)]
    t2 = time.perf_counter()
    w2.watch_json(iter(chunks), pod_from_line)
    t3 = time.perf_counter()
    b = len(sink.transport.bodies())
    print("%s: object loop %.3f s = %.0f events/s (%d notifications); raw loop (PodWatcher.watch_json) %.3f s = %.0f events/s (%d notifications)"
          % (envname, t1 - t0, n / (t1 - t0), a, t3 - t2, n / (t3 - t2), b))
