"""Device time of pte_run for the three filter regimes and both traces at a given size."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1_000_000_000
pods = int(float(sys.argv[2])) if len(sys.argv) > 2 else 10_000_000
from k8s_watcher_b200 import PhaseTransitionEngine
REG = {"staging": dict(ns_allow=None, critical_only=False), "development": dict(ns_allow=[0, 1], critical_only=False),
       "production": dict(ns_allow=[0, 1, 2, 3], critical_only=True)}
with PhaseTransitionEngine(n, pods) as e:
    e.generate(0x6B38735F77617463, n, pods, dist=0, p_del=0.05)
    for name, f in REG.items():
        e.set_filters(f["ns_allow"], critical_only=f["critical_only"])
        for dedupe in (True, False):
            best = min((e.run(dedupe=dedupe) for _ in range(3)), key=lambda s: s["ms_total"])
            print("%-12s dedupe=%d survived=%d emitted=%d ms_total=%.2f intern=%.2f sort=%.2f diff=%.2f assemble=%.2f" % (
                name, dedupe, best["n_survived"], best["n_emitted"], best["ms_total"], best["ms_intern"], best["ms_sort_key"],
                best["ms_diff"], best["ms_assemble"]), flush=True)
