"""Full-size parity: BASELINE configs 3 and 4 (1 B events / 10 M pods) -- counts and order-sensitive digest of the
emitted records against the C oracle's independent, strictly sequential generate+replay (oracle/oracle.c)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from k8s_watcher_b200 import PhaseTransitionEngine
from oracle import c_oracle
SEED = 0x6B38735F77617463
n, pods = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1_000_000_000, int(float(sys.argv[2])) if len(sys.argv) > 2 else 10_000_000
ok = True
with PhaseTransitionEngine(n, pods) as e:
    for name, kw in (("config 3 (uniform churn)", dict(dist=0, p_del=0.0)), ("config 4 (octave-Zipf + 5% delete draws)", dict(dist=1, p_del=0.05))):
        e.generate(SEED, n, pods, **kw)
        st = e.run()
        dig = e.digest()
        t0 = time.time()
        ost, odig = c_oracle.generate_replay(SEED, n, pods, **kw)
        same = all(st[k] == ost[k] for k in ost) and dig == odig
        ok &= same
        print("%s: n=%d pods=%d  engine emitted=%d tombstones=%d digest=%016x | oracle emitted=%d tombstones=%d digest=%016x (%.0f s CPU) -> %s"
              % (name, n, pods, st["n_emitted"], st["n_tombstones"], dig, ost["n_emitted"], ost["n_tombstones"], odig, time.time() - t0,
                 "MATCH" if same else "MISMATCH"), flush=True)
sys.exit(0 if ok else 1)
