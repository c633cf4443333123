"""Device-side speed of the raw watch-JSON path: line index + scanner kernel, CUDA-event times (no PCIe in them).
    python tools/json_bench.py [events per block, default 200000] [repeat, default 8]   -> about 1 GB of text"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import fake_kubernetes as fk
from k8s_watcher_b200 import PhaseTransitionEngine
from oracle import c_oracle
ne = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
rep = int(sys.argv[2]) if len(sys.argv) > 2 else 8
cols = c_oracle.generate(0x6B38735F77617463, ne, 3000, dist=1, p_del=0.05, p_nostatus=0.03, p_null=0.03)
block = fk.watch_json_lines(fk.events_from_columns(cols), seed=1)
if not block.endswith(b"\n"):
    block += b"\n"
text = block * rep
peak = 6546.9
with PhaseTransitionEngine(ne * rep + 16, 4000) as e:
    for _ in range(3):
        n = e.ingest_json(text)
        t = e.json_timing()
    gb = t["text_bytes"] / 1e9
    print("watch JSON: %d lines, %.3f GB of text (%.0f B/line): line index %.3f ms, scanner kernel %.3f ms = %.0f GB/s of text "
          "(%.3f of the measured %.1f GB/s HBM peak), %.1f M events/s on the device"
          % (n, gb, t["text_bytes"] / n, t["ms_index"], t["ms_parse"], gb / (t["ms_parse"] * 1e-3), gb / (t["ms_parse"] * 1e-3) / peak,
             peak, n / t["ms_parse"] / 1e3))

    # the same call from the host's point of view: where the text lives decides
    import time
    import numpy as np
    from k8s_watcher_b200.engine import PinnedBuffer
    pin = PinnedBuffer(len(text), np.uint8)
    pin.array[:] = np.frombuffer(text, np.uint8)
    for name, src in (("bytes object (pageable memory)", text), ("uint8 array in pinned memory", pin.array)):
        best = 1e9
        for _ in range(3):
            t0 = time.perf_counter(); e.ingest_json(src); best = min(best, time.perf_counter() - t0)
        print("  pte_ingest_json wall, text as %s: %.1f ms = %.1f GB/s of text, %.1f M events/s" % (name, best * 1e3, gb / best, n / best / 1e6))
    pin.free()

# pod objects of the size real clusters send (managedFields, several containers with env and mounts: 4-5 KB per event)
import json as _json
fat_meta = {"managedFields": [{"manager": "kube-controller-manager", "operation": "Update", "apiVersion": "v1", "time": "2025-01-01T00:00:00Z",
                               "fieldsType": "FieldsV1", "fieldsV1": {"f:metadata": {"f:labels": {".": {}, "f:app": {}, "f:pod-template-hash": {}},
                               "f:ownerReferences": {".": {}, "k:{\"uid\":\"00000000-dead-beef-0000-000000000000\"}": {}}},
                               "f:spec": {"f:containers": {"k:{\"name\":\"main\"}": {".": {}, "f:image": {}, "f:env": {}, "f:resources": {}}}}}}] * 3}
fat_spec = {"volumes": [{"name": "v%d" % k, "configMap": {"name": "cfg-%d" % k, "defaultMode": 420}} for k in range(4)],
            "containers2": [{"name": "c%d" % k, "image": "registry.local/team/service-%d:1.2.%d" % (k, k),
                             "env": [{"name": "ENV_%d" % j, "value": "value-%d" % j} for j in range(8)],
                             "volumeMounts": [{"name": "v%d" % j, "mountPath": "/etc/cfg/%d" % j, "readOnly": True} for j in range(4)],
                             "resources": {"limits": {"cpu": "500m", "memory": "512Mi"}, "requests": {"cpu": "100m", "memory": "128Mi"}}} for k in range(3)]}
extra_meta = _json.dumps(fat_meta, separators=(",", ":"))[1:-1].encode()
extra_spec = _json.dumps(fat_spec, separators=(",", ":"))[1:-1].encode()
fat_lines = []
for l in block.split(b"\n")[:40_000]:
    if not l:
        continue
    for marker, extra in ((b'"metadata":{', extra_meta), (b'"metadata": {', extra_meta), (b'"spec":{', extra_spec), (b'"spec": {', extra_spec)):
        i = l.find(marker)
        if i >= 0 and b'"object"' in l[:i]:
            l = l[:i + len(marker)] + extra + b"," + l[i + len(marker):]
    fat_lines.append(l)
fat = b"\n".join(fat_lines) + b"\n"
fat_text = fat * max(1, int(1.0e9 // len(fat)))
with PhaseTransitionEngine(len(fat_lines) * (len(fat_text) // len(fat)) + 16, 4000) as e:
    small = PhaseTransitionEngine(len(fat_lines) + 16, 4000)
    small.ingest_json(fat); fat_cols = small.read_log()
    small.ingest_json(b"\n".join(block.split(b"\n")[:40_000])); thin_cols = small.read_log(); small.close()
    assert all((fat_cols[k] == thin_cols[k]).all() for k in ("uid_hi", "uid_lo", "phase", "type", "ns_id")), "the padding changed what is read"
    for _ in range(3):
        n = e.ingest_json(fat_text)
        t = e.json_timing()
    gb = t["text_bytes"] / 1e9
    print("watch JSON, pods of realistic size: %d lines, %.3f GB of text (%.0f B/line): line index %.3f ms, scanner kernel %.3f ms = %.0f GB/s of text "
          "(%.3f of the measured %.1f GB/s HBM peak), %.1f M events/s on the device"
          % (n, gb, t["text_bytes"] / n, t["ms_index"], t["ms_parse"], gb / (t["ms_parse"] * 1e-3), gb / (t["ms_parse"] * 1e-3) / peak,
             peak, n / t["ms_parse"] / 1e3))

# request bodies built on the device (pte_serialize_payloads): device time of the kernel that writes them
import logging; logging.disable(logging.CRITICAL)
from k8s_watcher_b200 import PodWatcher, ReplayLog
from k8s_watcher_b200.pod_watcher import payload_fragments
os.chdir(ROOT)
w = PodWatcher("staging")
log = ReplayLog.synthetic(0x6B38735F77617463, 3_000_000, 40_000, p_del=0.05)
batch = w.replay(log)
frags, n_phase, n_ns = payload_fragments(w, log, "2025-06-01T12:00:00")
for _ in range(2):
    buf, off = w._engine.serialize_payloads(frags, n_phase, n_ns)
pt = w._engine.payload_timing()
print("payloads: %d request bodies, %.3f GB: payload_write_kernel %.3f ms = %.0f GB/s written (%.3f of the measured %.1f GB/s HBM peak)"
      % (len(batch), pt["body_bytes"] / 1e9, pt["ms_write"], pt["body_bytes"] / 1e9 / (pt["ms_write"] * 1e-3),
         pt["body_bytes"] / 1e9 / (pt["ms_write"] * 1e-3) / peak, peak))
