"""Sink side of the watcher: same surface as the reference's ClusterApiClient
(watcher/clusterapi_client.py:6-61), plus a batched call for the bulk engine.

The default transport is ``requests.Session`` like the reference's (:func:`requests_transport`).
``transport`` may be any callable ``(method, url, body_bytes_or_None, headers, timeout) -> status_code``;
the tests inject a recording one (tests/mock_transport.py), since there is no network where they run.
"""
from __future__ import annotations

import json
import logging
from typing import Any, Callable, Dict, List, Optional


class ConnectionFailed(Exception):
    """Raised by a transport that cannot reach the server."""


class RequestTimeout(Exception):
    """Raised by a transport whose request timed out."""


def requests_transport():
    """Transport over ``requests.Session`` (what the reference uses, clusterapi_client.py:9,36,58)."""
    import requests
    session = requests.Session()

    def send(method, url, body, headers, timeout):
        try:
            r = session.request(method, url, data=body, headers=headers, timeout=timeout)
        except requests.exceptions.ConnectionError as e:
            raise ConnectionFailed(str(e))
        except requests.exceptions.Timeout as e:
            raise RequestTimeout(str(e))
        return r.status_code
    return send


def wire_body(pod_data: Dict[str, Any]) -> bytes:
    """Bytes ``requests`` puts on the wire for ``session.post(endpoint, json=pod_data)``:
    ``json.dumps(pod_data, allow_nan=False)`` in insertion order, UTF-8 encoded."""
    return json.dumps(pod_data, allow_nan=False).encode("utf-8")


class ClusterApiClient:
    def __init__(self, base_url: str, api_key: Optional[str] = None, *, transport: Optional[Callable] = None,
                 timeout: Optional[float] = None, max_attempts: int = 1, delay_seconds: float = 0.0,
                 sleep: Callable[[float], None] = None):
        """``timeout``, ``max_attempts`` and ``delay_seconds`` are the knobs the reference's config declares
        (config/base.yaml: clusterapi.timeout, clusterapi.retry.*) but its code never reads; the defaults keep
        the reference's behaviour (one attempt, no timeout).  See :meth:`from_config`."""
        self.base_url = base_url.rstrip("/")
        self.api_key = api_key
        self.timeout = timeout
        self.max_attempts = max(1, int(max_attempts))
        self.delay_seconds = float(delay_seconds)
        import time as _time
        self._sleep = sleep if sleep is not None else _time.sleep
        self.transport = transport if transport is not None else requests_transport()
        self.session = self.transport            # the reference exposes .session; keep the name alive
        self.logger = logging.getLogger(__name__)
        # session.post(json=...) always sends this header (clusterapi_client.py:36); the bearer token only with a key (:14-18)
        self.headers: Dict[str, str] = {"Content-Type": "application/json"}
        if self.api_key:
            self.headers["Authorization"] = "Bearer %s" % self.api_key

    @classmethod
    def from_config(cls, config: Dict[str, Any], **kw) -> "ClusterApiClient":
        """Build the sink from the watcher's merged config the way the reference's (dead) factory meant to
        (pod_watcher.py:96-108): base_url, auth.api_key, timeout, plus the retry block nobody reads upstream."""
        c = config.get("clusterapi", {})
        retry = c.get("retry", {})
        return cls(c.get("base_url", "http://localhost:3000"), c.get("auth", {}).get("api_key") or None,
                   timeout=c.get("timeout", 30), max_attempts=retry.get("max_attempts", 1),
                   delay_seconds=retry.get("delay_seconds", 0), **kw)

    def _post_once(self, endpoint: str, body: bytes, name: str) -> bool:
        try:
            status = self.transport("POST", endpoint, body, self.headers, self.timeout)
            if status == 200:
                self.logger.info("Successfully updated pod data for %s", name)
                return True
            self.logger.error("Failed to update pod data. Status: %s", status)
            return False
        except ConnectionFailed:
            self.logger.error("Connection error: Unable to connect to clusterapi at %s", endpoint)
            return False
        except RequestTimeout:
            self.logger.error("Timeout error: Request to %s timed out", endpoint)
            return False
        except Exception as e:                    # clusterapi_client.py:51-53: never raise
            self.logger.error("Unexpected error calling clusterapi: %s", e)
            return False

    def _post_body(self, endpoint: str, body: bytes, name: str) -> bool:
        """One notification, up to ``max_attempts`` tries ``delay_seconds`` apart; False only if all fail."""
        for attempt in range(self.max_attempts):
            if self._post_once(endpoint, body, name):
                return True
            if attempt + 1 < self.max_attempts:
                self.logger.warning("Retrying %s (attempt %d of %d)", name, attempt + 2, self.max_attempts)
                if self.delay_seconds > 0:
                    self._sleep(self.delay_seconds)
        return False

    def update_pod_status(self, pod_data: Dict[str, Any]) -> bool:
        """POST one payload to ``{base}/api/pods/update``; True iff the answer is 200
        (clusterapi_client.py:20-53)."""
        endpoint = "%s/api/pods/update" % self.base_url
        self.logger.info("Calling clusterapi: %s", endpoint)
        try:
            body = wire_body(pod_data)
        except Exception as e:
            self.logger.error("Unexpected error calling clusterapi: %s", e)
            return False
        return self._post_body(endpoint, body, pod_data.get("name", "unknown"))

    def update_pod_status_batch(self, batch, event_timestamp: Optional[str] = None) -> int:
        """Hand every emitted record of ``batch`` to the sink, in replay order.
        Returns the number of successful updates."""
        endpoint = "%s/api/pods/update" % self.base_url
        ok = 0
        for body in batch.wire_bodies(event_timestamp=event_timestamp):      # serialised in bulk (on the GPU for column logs)
            ok += self._post_body(endpoint, body, "batch")
        return ok

    def health_check(self) -> bool:
        """GET ``{base}/health`` with a 5 s timeout (clusterapi_client.py:55-61)."""
        try:
            return self.transport("GET", "%s/health" % self.base_url, None, self.headers, 5) == 200
        except Exception:
            return False
