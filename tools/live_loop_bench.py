"""Per-batch cost of the live loop (PodWatcher(flush_every=4096) -> PTE_RUN_CARRY batches): wall time of
ingest + run + fetch for one 4096-event batch against the device time of the run, and the launch count.

    python tools/live_loop_bench.py [batch_events] [batches]
"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from k8s_watcher_b200 import PhaseTransitionEngine
from oracle import c_oracle

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
N = int(sys.argv[2]) if len(sys.argv) > 2 else 400
cols = c_oracle.generate(7, B * N, 20_000, dist=1, p_del=0.05)
with PhaseTransitionEngine(B, 1 << 20) as e:
    e.set_filters(None)
    walls, devs, launches = [], [], 0
    for k in range(N):
        part = {c: v[k * B:(k + 1) * B] for c, v in cols.items()}
        t0 = time.perf_counter()
        if os.environ.get("THREE_CALLS"):
            e.ingest(part)
            st = e.run(carry=True)
            got = e.fetch(columns=("idx", "type"))
        else:
            got, st = e.stream_batch(part)
        walls.append((time.perf_counter() - t0) * 1e6)
        devs.append(st["ms_total"] * 1e3)
        launches = st["n_launches"]
    w, d = np.array(walls[20:]), np.array(devs[20:])
    print("batch %d events, %d batches: wall per batch median %.1f us (p90 %.1f), device time of pte_run median %.1f us, "
          "%d kernel launches per run, wall/device %.2f, %.2f M events/s through the live loop"
          % (B, N, np.median(w), np.percentile(w, 90), np.median(d), launches, np.median(w) / np.median(d), B / np.median(w)))
