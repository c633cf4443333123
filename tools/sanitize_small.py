"""Small end-to-end run for compute-sanitizer (one tool per gpurun call): all kernels, both traces, filters, stream mode."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from k8s_watcher_b200 import PhaseTransitionEngine
from k8s_watcher_b200.engine import PinnedBuffer, partition_u64_static
from oracle import c_oracle
n, pods = 50_003, 700
with PhaseTransitionEngine(n, pods) as e:
    e.generate(7, n, pods, dist=1, p_del=0.05, p_nostatus=0.02, p_null=0.02)
    cols = e.read_log()
    for allow, crit in ((None, False), ([0, 1, 2, 3], True)):
        e.set_filters(allow, critical_only=crit)
        for dedupe in (True, False):
            st = e.run(dedupe=dedupe)
            got = e.fetch()
            want, ost, dig = c_oracle.replay(cols, critical_only=crit, ns_allow=allow, dedupe=dedupe, pods_hint=pods)
            assert np.array_equal(got.idx.astype(np.uint64), want) and e.digest() == dig
    e.set_filters(None)
    for lo, hi in ((0, 20_000), (20_000, n)):
        e.ingest({k: v[lo:hi] for k, v in cols.items()})
        e.run(carry=True)
    rec = np.random.default_rng(1).integers(0, 1 << 62, 30_001, dtype=np.uint64)
    assert np.array_equal(e.debug_sort(rec, 5, 41), rec[np.argsort((rec >> np.uint64(5)) & np.uint64((1 << 36) - 1), kind="stable")])
    a, b = PinnedBuffer(30_001, np.uint64), PinnedBuffer(30_001, np.uint64)
    a.array[:] = rec
    partition_u64_static(a.array, b.array, 8, 8)
    a.free(); b.free()
print("sanitize run ok")
