"""Development aid: the partition-first path against the C oracle at growing sizes, then timings at full size.

    PTE_PATH=partition python tools/try_partition.py [max_events]
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402

from k8s_watcher_b200 import PhaseTransitionEngine  # noqa: E402
from oracle import c_oracle  # noqa: E402

SEED = 0x6B38735F77617463
big = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1_000_000_000
FILTERS = {"staging": (None, False), "development": ([0, 1], False), "production": ([0, 1, 2, 3], True)}

ok = True
cases = [] if os.environ.get("SKIP_SMALL") else [(1000, 50, 0, 0.05), (100_000, 5_000, 0, 0.0), (300_000, 7_000, 1, 0.05), (3_000_000, 100_000, 0, 0.0),
         (10_000_000, 100_000, 0, 0.0), (20_000_000, 200_000, 1, 0.05), (5_000_000, 3, 1, 0.3)]
with PhaseTransitionEngine(20_000_000, 400_000) as e:
    for n, pods, dist, p_del in cases:
        cols = c_oracle.generate(SEED, n, pods, dist=dist, p_del=p_del, p_nostatus=0.01, p_null=0.01)
        for env, (allow, crit) in FILTERS.items():
            e.set_filters(allow, critical_only=crit)
            e.ingest(cols)
            st = e.run()
            got = e.fetch(columns=("idx",))
            idx, ost, dig = c_oracle.replay(cols, critical_only=crit, ns_allow=allow, pods_hint=pods)
            same = np.array_equal(got.idx.astype(np.uint64), idx) and all(st[k] == ost[k] for k in ost) and e.digest() == dig
            ok &= same
            print("n=%d pods=%d dist=%d %s: path=%d emitted=%d/%d pending=%d ms=%.3f -> %s" % (
                n, pods, dist, env, st["path"], st["n_emitted"], ost["n_emitted"], st["n_pending"], st["ms_total"],
                "MATCH" if same else "MISMATCH"), flush=True)
            if not same:
                a, b = got.idx.astype(np.uint64), idx
                m = min(len(a), len(b))
                d = np.nonzero(a[:m] != b[:m])[0]
                print("   first difference at", d[:5], a[d[:5]] if len(d) else None, b[d[:5]] if len(d) else None, len(a), len(b))

if big >= 100_000_000:
    try:
        with open(os.path.join(ROOT, "tests", "golden", "digests.json")) as f:
            rows = json.load(f)["rows"]
    except FileNotFoundError:
        rows = {"c3/1/0": {"digest": "b62ebd6c40660a15", "n_emitted": 802020055}, "c4/1/0": {"digest": "ed5bd1f7e0aee72d", "n_emitted": 820710277}}
    pods = big // 100
    with PhaseTransitionEngine(big, pods) as e:
        for name, kw in (("c3", dict(dist=0, p_del=0.0)), ("c4", dict(dist=1, p_del=0.05))):
            e.generate(SEED, big, pods, **kw)
            best = None
            for it in range(4):
                st = e.run()
                if best is None or st["ms_total"] < best["ms_total"]:
                    best = st
            dig = "%016x" % e.digest()
            want = rows.get("%s/1/0" % name) if big == 1_000_000_000 else None
            verdict = "unchecked" if want is None else ("MATCH" if (want["digest"] == dig and want["n_emitted"] == st["n_emitted"]) else "MISMATCH")
            ok &= verdict != "MISMATCH"
            print(name, {k: (round(v, 3) if isinstance(v, float) else v) for k, v in best.items()}, "digest", dig, verdict, flush=True)
sys.exit(0 if ok else 1)
