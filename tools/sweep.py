"""Tuning aid: per-stage device times of pte_run for each radix pass configuration."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1_000_000_000
pods = int(float(sys.argv[2])) if len(sys.argv) > 2 else 10_000_000
cfgs = sys.argv[3].split(",") if len(sys.argv) > 3 else ["0", "1", "2", "3", "4"]
dist = int(sys.argv[4]) if len(sys.argv) > 4 else 0
p_del = float(sys.argv[5]) if len(sys.argv) > 5 else 0.0
from k8s_watcher_b200 import PhaseTransitionEngine

for cfg in cfgs:
    os.environ["PTE_PASS_CFG"] = cfg
    with PhaseTransitionEngine(n, pods) as e:
        e.generate(0x6B38735F77617463, n, pods, dist=dist, p_del=p_del)
        best = None
        for it in range(4):
            st = e.run()
            if best is None or st["ms_total"] < best["ms_total"]:
                best = st
        print("cfg", cfg, {k: (round(v, 3) if isinstance(v, float) else v) for k, v in best.items()},
              "Gev/s %.2f" % (n / best["ms_total"] / 1e6), flush=True)
