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
