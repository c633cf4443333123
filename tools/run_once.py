"""Development aid: generate one workload on the device and replay it a few times (for ncu captures and knob sweeps).
    python tools/run_once.py [events] [pods] [runs] [dist] [p_del]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from k8s_watcher_b200 import PhaseTransitionEngine
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1_000_000_000
pods = int(float(sys.argv[2])) if len(sys.argv) > 2 else n // 100
runs = int(sys.argv[3]) if len(sys.argv) > 3 else 3
dist = int(sys.argv[4]) if len(sys.argv) > 4 else 0
p_del = float(sys.argv[5]) if len(sys.argv) > 5 else 0.0
with PhaseTransitionEngine(n, pods) as e:
    e.generate(0x6B38735F77617463, n, pods, dist=dist, p_del=p_del)
    for _ in range(runs):
        st = e.run()
        print({k: (round(v, 3) if isinstance(v, float) else v) for k, v in st.items()}, "digest %016x" % e.digest(), flush=True)
