"""In-process stand-in for the ``kubernetes`` package (not installed here, no network).

Provides exactly the names the reference touches
(/root/reference/watcher/pod_watcher.py:1-2,16,118,129,134,137,140,146,264,277):
``client.CoreV1Api``, ``config.load_kube_config``, ``config.load_incluster_config``,
``config.ConfigException``, ``watch.Watch().stream/stop``.  ``install()`` puts it in
``sys.modules`` so both the real reference (golden generation, this container only)
and the drop-in ``PodWatcher`` can ``from kubernetes import client, config, watch``.

Also holds the V1Pod-like value classes and the deterministic pod factory used
by the fixtures.
"""
from __future__ import annotations

import datetime as _dt
import sys
import types
from typing import Any, Dict, Iterable, List, Optional

# events the next CoreV1Api().list_pod_for_all_namespaces "stream" will yield
_PENDING_STREAM: List[Dict[str, Any]] = []


class ConfigException(Exception):
    pass


class _Obj:
    def __init__(self, **kw):
        self.__dict__.update(kw)

    def __repr__(self):  # str(cs.state) goes into the payload (pod_watcher.py:181)
        return "%s(%s)" % (type(self).__name__, ", ".join("%s=%r" % kv for kv in sorted(self.__dict__.items())))


class V1ObjectMeta(_Obj):
    pass


class V1PodStatus(_Obj):
    pass


class V1PodSpec(_Obj):
    pass


class V1PodCondition(_Obj):
    pass


class V1ContainerStatus(_Obj):
    pass


class V1ContainerState(_Obj):
    pass


class V1Container(_Obj):
    pass


class V1Pod(_Obj):
    pass


class CoreV1Api:
    def get_api_version(self):
        return "v1"

    def list_namespace(self, **kw):
        return _Obj(items=[_Obj(metadata=_Obj(name=n)) for n in ("default", "kube-system")])

    def list_pod_for_all_namespaces(self, **kw):
        return _Obj(items=[e["object"] for e in _PENDING_STREAM])


class Watch:
    def __init__(self):
        self.stopped = False

    def stream(self, func, *a, **kw):
        for ev in list(_PENDING_STREAM):
            if self.stopped:
                return
            yield ev

    def stop(self):
        self.stopped = True


def set_stream(events: Iterable[Dict[str, Any]]) -> None:
    _PENDING_STREAM[:] = list(events)


def install() -> types.ModuleType:
    k = types.ModuleType("kubernetes")
    client = types.ModuleType("kubernetes.client")
    client.CoreV1Api = CoreV1Api
    config = types.ModuleType("kubernetes.config")
    config.ConfigException = ConfigException
    config.load_kube_config = lambda *a, **kw: None
    config.load_incluster_config = lambda *a, **kw: None
    watch = types.ModuleType("kubernetes.watch")
    watch.Watch = Watch
    k.client, k.config, k.watch = client, config, watch
    sys.modules["kubernetes"] = k
    sys.modules["kubernetes.client"] = client
    sys.modules["kubernetes.config"] = config
    sys.modules["kubernetes.watch"] = watch
    return k


# ---------------------------------------------------------------------------
# deterministic pod factory (fixture contract)
# ---------------------------------------------------------------------------
NAMESPACES = ("default", "kube-system", "production", "monitoring", "other")
_T0 = _dt.datetime(2025, 1, 1, 0, 0, 0)


def make_pod(p: int, uid: str, namespace: str, phase: Optional[str], *, has_status: bool = True,
             rich: bool = True, resource_version: Optional[int] = None) -> V1Pod:
    """V1Pod-like object for synthetic pod ``p``.

    ``rich`` pods carry labels, one container, one condition and one container
    status so every list comprehension of _extract_pod_data (:168-192) runs;
    every 7th pod has ``spec=None`` and every 5th ``labels=None`` /
    ``creation_timestamp=None`` to hit the ``or {}`` / ``else None`` arms.
    """
    md = V1ObjectMeta(
        name="pod-%d" % p, namespace=namespace, uid=uid,
        labels=None if p % 5 == 4 else ({"app": "svc-%d" % (p % 9), "tier": "t%d" % (p % 3)} if rich else {}),
        annotations=None if p % 3 == 2 else ({"owner": "team-%d" % (p % 4)} if rich else {}),
        creation_timestamp=None if p % 5 == 3 else _T0 + _dt.timedelta(seconds=p),
        resource_version=None if resource_version is None else str(resource_version),
    )
    status = None
    if has_status:
        conds = [V1PodCondition(type="Ready", status="True" if phase == "Running" else "False",
                                reason=None, message=None)] if rich and p % 2 == 0 else None
        cstats = [V1ContainerStatus(name="main", ready=phase == "Running", restart_count=p % 4,
                                    state=V1ContainerState(running=phase == "Running") if p % 4 else None)] \
            if rich and p % 3 != 1 else None
        status = V1PodStatus(phase=phase, conditions=conds, container_statuses=cstats)
    spec = None if p % 7 == 6 else V1PodSpec(
        node_name="node-%d" % (p % 11) if p % 6 else None,
        containers=[V1Container(name="main", image="registry.local/app:%d" % (p % 13))] if rich else None)
    return V1Pod(metadata=md, status=status, spec=spec)


def uid_str(hi: int, lo: int) -> str:
    """128-bit uid as the canonical 8-4-4-4-12 hex form k8s uses for metadata.uid."""
    h = "%016x%016x" % (hi, lo)
    return "-".join((h[0:8], h[8:12], h[12:16], h[16:20], h[20:32]))


PHASE_STR = ("Pending", "Running", "Succeeded", "Failed", "Unknown")
TYPE_STR = ("ADDED", "MODIFIED", "DELETED")


def events_from_columns(cols, *, rich: bool = True):
    """Column log (oracle/c_oracle.generate layout) -> list of watch-event dicts
    ``{'type': str, 'object': V1Pod}`` in arrival order (pod_watcher.py:267-268)."""
    out = []
    n_shape = len(cols["rv"])
    for i in range(n_shape):
        hi, lo = int(cols["uid_hi"][i]), int(cols["uid_lo"][i])
        code = int(cols["phase"][i])
        has_status = code != 6
        phase = PHASE_STR[code] if code < 5 else None
        p = int(cols["pod"][i]) if "pod" in cols else (hi % 1000003)
        pod = make_pod(p, uid_str(hi, lo), NAMESPACES[int(cols["ns_id"][i])], phase, has_status=has_status,
                       rich=rich, resource_version=int(cols["rv"][i]))
        out.append({"type": TYPE_STR[int(cols["type"][i])], "object": pod})
    return out


def watch_json_lines(events, *, seed: int = 0, compact: bool = True) -> bytes:
    """The watch stream as the Kubernetes API would send it: one JSON watch event per line
    (``{"type": ..., "object": {...pod...}}``, API field names), keys in a per-line shuffled order, plus
    look-alike keys deeper down (ownerReferences[].uid, an annotation whose text contains ``"phase"``) that a
    parser must not confuse with the fields the watcher reads (pod_watcher.py:162-167, :216-217)."""
    import json
    import random
    rnd = random.Random(seed)

    def shuffled(d):
        items = list(d.items())
        rnd.shuffle(items)
        return dict(items)

    lines = []
    for ev in events:
        pod = ev["object"]
        md = pod.metadata
        meta = {"name": md.name, "namespace": md.namespace, "uid": md.uid,
                "ownerReferences": [{"kind": "ReplicaSet", "uid": "00000000-dead-beef-0000-000000000000", "name": "rs"}],
                "annotations": dict(md.annotations or {}, note='fake {"phase":"Failed","uid":"x"} \\ "quoted"'),
                "labels": md.labels or {}}
        if md.resource_version is not None:
            meta["resourceVersion"] = md.resource_version
        if md.creation_timestamp is not None:
            meta["creationTimestamp"] = md.creation_timestamp.isoformat() + "Z"
        obj = {"kind": "Pod", "apiVersion": "v1", "metadata": shuffled(meta)}
        if pod.spec is not None:
            obj["spec"] = {"nodeName": pod.spec.node_name,
                           "containers": [{"name": c.name, "image": c.image, "env": [{"name": "phase", "value": "Running"}]}
                                          for c in (pod.spec.containers or [])]}
        if pod.status is not None:
            st = {"conditions": [{"type": c.type, "status": c.status, "phase": "decoy"} for c in (pod.status.conditions or [])],
                  "hostIP": "10.0.0.1"}
            if pod.status.phase is not None or rnd.random() < 0.5:
                st["phase"] = pod.status.phase            # None -> JSON null; sometimes the key is simply absent
            obj["status"] = shuffled(st)
        elif rnd.random() < 0.5:
            obj["status"] = None                           # "status": null is as falsy as a missing status
        line = json.dumps(shuffled({"type": ev["type"], "object": shuffled(obj)}),
                          separators=(",", ":") if (compact and rnd.random() < 0.7) else None)
        lines.append(line)
    return ("\n".join(lines) + ("\n" if rnd.random() < 0.5 or not lines else "")).encode()
