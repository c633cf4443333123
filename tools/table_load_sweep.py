"""How does intern_pack's time depend on the dictionary's size and load?  (same 1e9 events, pods and table size varied)
    python tools/table_load_sweep.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from k8s_watcher_b200 import PhaseTransitionEngine, PteError
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1_000_000_000
for max_pods, pods_list in ((10_000_000, (3_000_000, 5_000_000, 7_000_000, 10_000_000, 12_000_000, 14_000_000, 15_500_000)),
                            (5_000_000, (3_000_000, 5_000_000, 6_000_000, 7_000_000, 7_800_000)),
                            (2_500_000, (2_500_000, 3_500_000))):
    with PhaseTransitionEngine(n, max_pods) as e:
        for pods in pods_list:
            try:
                e.generate(0x6B38735F77617463, n, pods)
                best = min((e.run() for _ in range(3)), key=lambda s: s["ms_intern"])
                slots = 256
                while slots < max_pods / 0.6 + 1:
                    slots *= 2
                print("table %4d MB  pods %9d  load %.2f  intern %.2f ms  total %.2f ms" % (slots * 16 >> 20, pods, pods / slots, best["ms_intern"], best["ms_total"]), flush=True)
            except PteError as ex:
                print("table for max_pods %d, pods %d: %s" % (max_pods, pods, ex), flush=True)
