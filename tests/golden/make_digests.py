#!/usr/bin/env python
"""Writes tests/golden/digests.json: what the C oracle (oracle/oracle.c: orc_generate_replay, the strictly
sequential restatement of SURVEY.md 8a row 7 over its own generator) says about every full-size workload that
bench.py and the -m gpu tests replay -- emitted count, tombstones and the order-sensitive digest of the
emitted records, per uid-hash shard g of G.

    python tests/golden/make_digests.py [--jobs 8]

Rows are keyed "workload/G/g".  BASELINE config 5 (8 B events / 80 M pods over 2/4/8 GPUs) uses the G = 8 rows
at every GPU count: with 2 or 4 GPUs each rank replays 4 or 2 of the eight 1 B-event shards one after the other.
About 30 s of one core per 1 B-event row.
"""
import argparse
import json
import os
import sys
from concurrent.futures import ProcessPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

SEED = 0x6B38735F77617463
# name: (events per shard, pods per shard, dist, p_del)   -- the same table as bench.py WORKLOADS
WORKLOADS = {
    "c2": (10_000_000, 100_000, 0, 0.0),
    "c3": (1_000_000_000, 10_000_000, 0, 0.0),
    "c4": (1_000_000_000, 10_000_000, 1, 0.05),
}
SHARDINGS = {"c2": (1, 2), "c3": (1, 2, 4, 8), "c4": (1, 2)}


def one(job):
    name, G, g = job
    from oracle import c_oracle
    n, pods, dist, p_del = WORKLOADS[name]
    st, dig = c_oracle.generate_replay(SEED, n, pods, dist=dist, p_del=p_del, shard=g, n_shards=G)
    return "%s/%d/%d" % (name, G, g), {"n_in": st["n_in"], "n_survived": st["n_survived"], "n_emitted": st["n_emitted"],
                                       "n_tombstones": st["n_tombstones"], "digest": "%016x" % dig}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--jobs", type=int, default=os.cpu_count() or 1)
    ap.add_argument("--out", default=os.path.join(ROOT, "tests", "golden", "digests.json"))
    args = ap.parse_args()
    jobs = [(name, G, g) for name, gs in SHARDINGS.items() for G in gs for g in range(G)]
    rows = {}
    with ProcessPoolExecutor(args.jobs) as ex:
        for key, row in ex.map(one, jobs):
            rows[key] = row
            print(key, row, flush=True)
    doc = {"seed": hex(SEED), "generator": "oracle/oracle.c orc_generate_replay (dedupe on, staging filters)",
           "workloads": {k: dict(zip(("events_per_shard", "pods_per_shard", "dist", "p_del"), v)) for k, v in WORKLOADS.items()},
           "rows": dict(sorted(rows.items()))}
    with open(args.out, "w") as f:
        json.dump(doc, f, indent=1)
        f.write("\n")


if __name__ == "__main__":
    main()
