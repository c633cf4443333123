"""Minimal stand-ins for the ``kubernetes`` names the watcher uses, for hosts where the
package is not installed (this image has no network).  A real ``kubernetes`` install wins.

``CoreV1Api.list_pod_for_all_namespaces`` / ``Watch.stream`` replay whatever events were
queued with :func:`queue_events`, so the live-loop surface can be driven from a replay log.
"""
from __future__ import annotations

import types
from typing import Any, Dict, Iterable, List

_QUEUE: List[Dict[str, Any]] = []


def queue_events(events: Iterable[Dict[str, Any]]) -> None:
    _QUEUE[:] = list(events)


class ConfigException(Exception):
    pass


class _NS:
    def __init__(self, **kw):
        self.__dict__.update(kw)


class CoreV1Api:
    def get_api_version(self):
        return "v1 (replay stub)"

    def list_namespace(self, **kw):
        names = sorted({e["object"].metadata.namespace for e in _QUEUE})
        return _NS(items=[_NS(metadata=_NS(name=n)) for n in names])

    def list_pod_for_all_namespaces(self, **kw):
        return _NS(items=[e["object"] for e in _QUEUE])


class Watch:
    def __init__(self):
        self._stop = False

    def stream(self, func, *args, **kwargs):
        for ev in list(_QUEUE):
            if self._stop:
                break
            yield ev

    def stop(self):
        self._stop = True


client = types.SimpleNamespace(CoreV1Api=CoreV1Api)
config = types.SimpleNamespace(
    ConfigException=ConfigException,
    load_kube_config=lambda *a, **kw: None,
    load_incluster_config=lambda *a, **kw: None,
)
watch = types.SimpleNamespace(Watch=Watch)
