"""Measurements for the rows of SURVEY 8f (stream batches, raw watch JSON, payload serialisation, R-trace, filters)."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import logging; logging.disable(logging.CRITICAL)
import numpy as np
import fake_kubernetes as fk
from k8s_watcher_b200 import PhaseTransitionEngine, PodWatcher, ReplayLog
from k8s_watcher_b200.pod_watcher import payload_fragments
from oracle import c_oracle
os.chdir(ROOT)
SEED = 0x6B38735F77617463

# 1. stream: 200 M events in 20 carried batches of 10 M, device-resident generation per batch is not possible
#    (one log per engine), so batches are ingested from pinned host memory; device time only.
n, pods, B = 200_000_000, 2_000_000, 10_000_000
with PhaseTransitionEngine(n, pods) as e:
    e.generate(SEED, n, pods, p_del=0.02)
    st_all = e.run()
    cols = e.read_log()
with PhaseTransitionEngine(B, pods) as e:
    ms, emitted = 0.0, 0
    for lo in range(0, n, B):
        e.ingest({k: v[lo:lo + B] for k, v in cols.items()})
        st = e.run(carry=True)
        ms += st["ms_total"]; emitted += st["n_emitted"]
    assert emitted == st_all["n_emitted"]
    print("stream: %d events in %d carried batches of %d: %.1f ms device time (%.2f G events/s), emitted %d == one replay (%.1f ms)"
          % (n, n // B, B, ms, n / ms / 1e6, emitted, st_all["ms_total"]))
del cols

# 2. raw watch JSON
ne, pj = 200_000, 3_000
events = fk.events_from_columns(c_oracle.generate(SEED, ne, pj, dist=1, p_del=0.05, p_nostatus=0.03, p_null=0.03))
text = fk.watch_json_lines(events, seed=1)
with PhaseTransitionEngine(ne, pj) as e:
    e.ingest_json(text)
    t0 = time.perf_counter()
    for _ in range(3):
        e.ingest_json(text)
    dt = (time.perf_counter() - t0) / 3
    print("watch JSON: %d events, %.1f MB of text: pte_ingest_json %.1f ms wall incl. H2D (%.2f GB/s, %.2f M events/s)"
          % (ne, len(text) / 1e6, dt * 1e3, len(text) / dt / 1e9, ne / dt / 1e6))
    t0 = time.perf_counter()
    for line in text.split(b"\n")[:20000]:
        if line: json.loads(line)
    print("            CPython json.loads alone: %.2f M events/s on one core" % (20000 / (time.perf_counter() - t0) / 1e6))

# 3. payload serialisation
w = PodWatcher("staging")
log = ReplayLog.synthetic(SEED, 4_000_000, 40_000, p_del=0.05)
batch = w.replay(log)
frags, n_phase, n_ns = payload_fragments(w, log, "2025-06-01T12:00:00")
eng = w._engine
eng.serialize_payloads(frags, n_phase, n_ns, count=1000)
t0 = time.perf_counter()
buf, off = eng.serialize_payloads(frags, n_phase, n_ns)
dt = time.perf_counter() - t0
print("payloads: %d request bodies, %.1f MB: pte_serialize_payloads %.1f ms wall incl. D2H (%.2f GB/s, %.1f M bodies/s)"
      % (len(batch), buf.size / 1e6, dt * 1e3, buf.size / dt / 1e9, len(batch) / dt / 1e6))
t0 = time.perf_counter()
k = 0
for p in batch.payloads(event_timestamp="2025-06-01T12:00:00"):
    json.dumps(p, allow_nan=False); k += 1
    if k == 50_000: break
print("          CPython dict + json.dumps: %.3f M bodies/s on one core" % (k / (time.perf_counter() - t0) / 1e6))
