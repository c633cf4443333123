"""CPU oracle for the bulk phase-transition path.  TEST INFRASTRUCTURE ONLY.

Nothing under ``k8s_watcher_b200/`` may import this module; only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs may.  It restates, in plain Python, what the reference watcher does to one
pod-watch event, plus the last-known-phase layer that BASELINE.json's north_star
asks for (which the reference does not implement).

Parity status
-------------
* Filters + payload (the "R-trace"): PINNED.  ``tests/golden/make_golden.py``
  runs the real, unmodified ``/root/reference/watcher/pod_watcher.py`` under a
  stub ``kubernetes`` module and commits what it produced to ``tests/golden``;
  ``tests/test_oracle_golden.py`` checks this restatement against those files.
* Transition layer (the "T-trace"): PARITY UNPINNED.  The reference holds no
  per-pod state (pod_watcher.py:214-241 is a stateless filter and the POST at
  :236 is commented out), so nothing upstream can pin it.  It is pinned only by
  the sequential spec in SURVEY.md section 8a row 7, restated in
  :func:`transition_trace` below.

Reference lines followed (all in /root/reference/watcher/pod_watcher.py):
  :204-212  should_process_event   -> :func:`should_process_event`
  :214-233  handle_pod_event       -> :func:`reaches_post_site`
  :159-202  _extract_pod_data      -> :func:`extract_pod_data`
  :47-75    config merge / env substitution -> :func:`merge_configs`,
            :func:`substitute_env_vars`
"""
from __future__ import annotations

import os
from typing import Any, Dict, Iterable, List, Optional, Sequence, Tuple

# ---------------------------------------------------------------------------
# Integer codes shared by the replay-log format (kept independent from the
# product's copy in k8s_watcher_b200/codes.py; tests assert they agree).
# ---------------------------------------------------------------------------
TYPE_NAMES = ("ADDED", "MODIFIED", "DELETED")
TYPE_ADDED, TYPE_MODIFIED, TYPE_DELETED = 0, 1, 2

PHASE_NAMES = ("Pending", "Running", "Succeeded", "Failed", "Unknown")
PH_PENDING, PH_RUNNING, PH_SUCCEEDED, PH_FAILED, PH_UNKNOWN = 0, 1, 2, 3, 4
PH_NULL = 5        # status object present, status.phase is None
PH_NOSTATUS = 6    # pod.status is None  -> payload phase "Unknown" (pod_watcher.py:167)
PH_FIRST_DYNAMIC = 7  # codes >= 7: other phase strings interned by the log


def effective_phase(code: int) -> int:
    """Phase the payload would carry (pod_watcher.py:167): no status -> "Unknown"."""
    return PH_UNKNOWN if code == PH_NOSTATUS else code


# ---------------------------------------------------------------------------
# config handling (pod_watcher.py:47-75), restated
# ---------------------------------------------------------------------------
def merge_configs(base: Dict[str, Any], override: Dict[str, Any]) -> Dict[str, Any]:
    """Recursive dict merge, override wins, lists replaced wholesale (:47-57)."""
    out = dict(base)
    for k, v in override.items():
        if isinstance(out.get(k), dict) and isinstance(v, dict):
            out[k] = merge_configs(out[k], v)
        else:
            out[k] = v
    return out


def substitute_env_vars(obj: Any) -> Any:
    """``${VAR}`` / ``${VAR:-default}`` only when the whole string is the placeholder (:59-75)."""
    if isinstance(obj, dict):
        return {k: substitute_env_vars(v) for k, v in obj.items()}
    if isinstance(obj, list):
        return [substitute_env_vars(v) for v in obj]
    if isinstance(obj, str) and obj.startswith("${") and obj.endswith("}"):
        body, default = obj[2:-1], ""
        if ":-" in body:
            body, default = body.split(":-", 1)
        return os.getenv(body, default)
    return obj


# ---------------------------------------------------------------------------
# per-event logic on pod OBJECTS (what the reference sees)
# ---------------------------------------------------------------------------
def should_process_event(environment: str, config: Dict[str, Any], event_type: str, pod) -> bool:
    """pod_watcher.py:204-212."""
    if environment == "production":
        critical_only = config.get("watcher", {}).get("alerts", {}).get("critical_events_only", False)
        if (critical_only and event_type not in ["DELETED"] and pod.status
                and pod.status.phase not in ["Failed", "Succeeded"]):
            return False
    return True


def reaches_post_site(environment: str, config: Dict[str, Any], event_type: str, pod) -> bool:
    """True iff handle_pod_event would get to the (commented-out) POST at :236.

    Order of checks as in pod_watcher.py:220 then :226-229.
    """
    if not should_process_event(environment, config, event_type, pod):
        return False
    target = config.get("watcher", {}).get("namespaces", [])
    if target and pod.metadata.namespace not in target:
        return False
    return True


def extract_pod_data(environment: str, pod, event_timestamp: str) -> Dict[str, Any]:
    """pod_watcher.py:159-202 with ``datetime.now().isoformat()`` (:199) injected.

    Key order is the reference's insertion order; the caller appends
    ``event_type`` last (:233).
    """
    st, sp, md = pod.status, pod.spec, pod.metadata
    return {
        "name": md.name,
        "namespace": md.namespace,
        "uid": md.uid,
        "environment": environment,
        "status": {
            "phase": st.phase if st else "Unknown",
            "conditions": [
                {"type": c.type, "status": c.status, "reason": c.reason, "message": c.message}
                for c in (st.conditions or [])
            ] if st else [],
            "container_statuses": [
                {"name": cs.name, "ready": cs.ready, "restart_count": cs.restart_count,
                 "state": str(cs.state) if cs.state else None}
                for cs in (st.container_statuses or [])
            ] if st else [],
        },
        "spec": {
            "node_name": sp.node_name if sp else None,
            "containers": [{"name": c.name, "image": c.image} for c in (sp.containers or [])] if sp else [],
        },
        "metadata": {
            "labels": md.labels or {},
            "annotations": md.annotations or {},
            "creation_timestamp": md.creation_timestamp.isoformat() if md.creation_timestamp else None,
        },
        "event_timestamp": event_timestamp,
    }


def reference_trace(environment: str, config: Dict[str, Any], events: Iterable[Tuple[str, Any]],
                    event_timestamp: str = "1970-01-01T00:00:00") -> List[Dict[str, Any]]:
    """R-trace: payloads of all events that reach the POST site, in arrival order."""
    out = []
    for event_type, pod in events:
        if reaches_post_site(environment, config, event_type, pod):
            d = extract_pod_data(environment, pod, event_timestamp)
            d["event_type"] = event_type
            out.append(d)
    return out


def transition_filter(payloads: Sequence[Dict[str, Any]]) -> List[int]:
    """T-trace layer over an R-trace (no reference counterpart).

    Returns positions (into ``payloads``) that a last-known-phase watcher would
    notify: every DELETED (and forget the pod), otherwise only when the payload
    phase differs from the last one notified for that uid.
    """
    last: Dict[str, Any] = {}
    keep = []
    for i, p in enumerate(payloads):
        uid, phase = p["uid"], p["status"]["phase"]
        if p["event_type"] == "DELETED":
            keep.append(i)
            last.pop(uid, None)
        elif uid not in last or last[uid] != phase:
            keep.append(i)
            last[uid] = phase
    return keep


# ---------------------------------------------------------------------------
# the same logic on integer replay columns (what the engine sees)
# ---------------------------------------------------------------------------
def survives_filters(type_code: int, phase_code: int, ns_id: int,
                     critical_only: bool, ns_allow: Optional[frozenset]) -> bool:
    """Column form of :func:`reaches_post_site`.

    ``critical_only`` is ``environment == 'production' and critical_events_only``.
    ``pod.status`` truthy <=> phase_code != PH_NOSTATUS; a null phase is "not in
    ['Failed','Succeeded']" so it is dropped (:209).  ``ns_allow`` None/empty
    means no namespace filter (:227).
    """
    if critical_only and type_code != TYPE_DELETED and phase_code != PH_NOSTATUS \
            and phase_code not in (PH_FAILED, PH_SUCCEEDED):
        return False
    if ns_allow and ns_id not in ns_allow:
        return False
    return True


def replay_columns(uid_hi, uid_lo, rv, ns_id, phase, type_, *, critical_only: bool = False,
                   ns_allow: Optional[Iterable[int]] = None, dedupe: bool = True) -> Tuple[List[int], Dict[str, int]]:
    """Sequential replay of a column log.  Returns (emitted arrival indices, stats).

    Events are replayed in ascending (rv, arrival index) order; for an ordered
    log (rv non-decreasing) that is arrival order.  ``dedupe=False`` gives the
    R-trace, ``dedupe=True`` the T-trace.
    """
    n = len(rv)
    allow = frozenset(int(x) for x in ns_allow) if ns_allow else None
    order = range(n)
    if any(rv[i] > rv[i + 1] for i in range(n - 1)):
        order = sorted(range(n), key=lambda i: (int(rv[i]), i))
    last: Dict[Tuple[int, int], int] = {}
    out: List[int] = []
    survived = tomb = 0
    for i in order:
        t, ph = int(type_[i]), int(phase[i])
        if not survives_filters(t, ph, int(ns_id[i]), critical_only, allow):
            continue
        survived += 1
        if not dedupe:
            out.append(i)
            tomb += t == TYPE_DELETED
            continue
        key = (int(uid_hi[i]), int(uid_lo[i]))
        eff = effective_phase(ph)
        if t == TYPE_DELETED:
            out.append(i)
            tomb += 1
            last.pop(key, None)
        elif key not in last or last[key] != eff:
            out.append(i)
            last[key] = eff
    return out, {"n_in": n, "n_survived": survived, "n_emitted": len(out), "n_tombstones": tomb}


# ---------------------------------------------------------------------------
# order-sensitive digest of an emitted record list (spec shared with the engine
# and with oracle/oracle.c; all arithmetic mod 2**64)
# ---------------------------------------------------------------------------
_M = (1 << 64) - 1


def mix64(z: int) -> int:
    z &= _M
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _M
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _M
    return z ^ (z >> 31)


def record_digest(position: int, rv: int, uid_hi: int, uid_lo: int, ns_id: int, phase: int, type_: int) -> int:
    small = ((ns_id & 0xFFFFFFFF) << 16) | ((phase & 0xFF) << 8) | (type_ & 0xFF)
    h = mix64(uid_lo ^ small)
    h = mix64(uid_hi ^ h)
    h = mix64(rv ^ h)
    return mix64(((position + 1) * 0x9E3779B97F4A7C15 & _M) ^ h)


def digest(emitted: Sequence[int], uid_hi, uid_lo, rv, ns_id, phase, type_) -> int:
    d = 0
    for pos, i in enumerate(emitted):
        d = (d + record_digest(pos, int(rv[i]), int(uid_hi[i]), int(uid_lo[i]),
                               int(ns_id[i]), int(phase[i]), int(type_[i]))) & _M
    return d
