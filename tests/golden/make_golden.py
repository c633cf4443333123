#!/usr/bin/env python
"""Generate tests/golden/*.json.gz by running the REAL, unmodified reference watcher.

Runs only in the build container (needs /root/reference, which does not travel
to the GPU box).  Usage:  python tests/golden/make_golden.py

For each environment (development / staging / production) and each stream it
imports /root/reference/watcher/pod_watcher.py under the stub ``kubernetes``
package (tests/fake_kubernetes.py), drives the stream through
``PodWatcher.handle_pod_event`` AND through ``PodWatcher.start_watching`` (the
live-loop surface, pod_watcher.py:243-277), and records every payload that
reaches the POST site (:232-236).  ``datetime.now()`` (:199) is frozen.

Streams (input columns come from oracle/liboracle.so's sequential generator):
  c1     BASELINE config 1: 1 000 events / 50 pods, uniform churn, 5 % delete draws
  edge   400 events / 12 pods with status=None and phase=None pods mixed in
Fixture layout: {"env", "config", "frozen_timestamp", "stream", "gen": {...},
 "columns": {...lists...}, "r_trace": [arrival indices], "payloads": [...]}.
"""
from __future__ import annotations

import datetime as real_datetime
import gzip
import json
import logging
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.dont_write_bytecode = True

import fake_kubernetes as fk  # noqa: E402
from oracle import c_oracle  # noqa: E402

FROZEN = "2025-06-01T12:00:00"
SEED = 0x6B38735F77617463

STREAMS = {
    "c1": dict(n=1000, pods=50, dist=0, p_del=0.05, p_nostatus=0.0, p_null=0.0),
    "edge": dict(n=400, pods=12, dist=1, p_del=0.10, p_nostatus=0.08, p_null=0.08),
}


class _FrozenDatetime(real_datetime.datetime):
    @classmethod
    def now(cls, tz=None):
        return real_datetime.datetime.fromisoformat(FROZEN)


def main() -> None:
    fk.install()
    os.chdir(REF)                      # config paths are CWD-relative (pod_watcher.py:22,25)
    sys.path.insert(0, REF)
    import watcher.pod_watcher as ref  # the real thing
    ref.datetime = _FrozenDatetime
    logging.disable(logging.CRITICAL)

    for sname, sp in STREAMS.items():
        cols = c_oracle.generate(SEED, sp["n"], sp["pods"], dist=sp["dist"], p_del=sp["p_del"],
                                 p_nostatus=sp["p_nostatus"], p_null=sp["p_null"])
        events = fk.events_from_columns(cols)
        for env in ("development", "staging", "production"):
            w = ref.PodWatcher(environment=env)
            recorded = []
            orig = w._extract_pod_data

            def rec(pod, _orig=orig, _r=recorded):
                d = _orig(pod)
                _r.append(d)          # handle_pod_event adds event_type to this same dict (:233)
                return d

            w._extract_pod_data = rec
            arrival = []
            for i, ev in enumerate(events):
                before = len(recorded)
                w.handle_pod_event(ev["type"], ev["object"])
                if len(recorded) > before:
                    arrival.append(i)
            direct = [dict(d) for d in recorded]

            # live-loop surface: same events through start_watching()
            recorded.clear()
            fk.set_stream(events)
            w.start_watching()
            assert [dict(d) for d in recorded] == direct, "start_watching and handle_pod_event disagree"

            fixture = {
                "env": env, "config": w.config, "frozen_timestamp": FROZEN, "stream": sname,
                "gen": dict(seed=SEED, **sp),
                "columns": {k: [int(x) for x in v] for k, v in cols.items()},
                "r_trace": arrival, "payloads": direct,
            }
            out = os.path.join(HERE, "ref_%s_%s.json.gz" % (sname, env))
            with gzip.GzipFile(out, "wb", mtime=0) as f:
                f.write(json.dumps(fixture, sort_keys=False, separators=(",", ":")).encode())
            print("%s: %d events -> %d reach the POST site (%d bytes)" % (
                os.path.basename(out), len(events), len(direct), os.path.getsize(out)))


if __name__ == "__main__":
    main()
