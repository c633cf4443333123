"""Recording transport for ClusterApiClient tests (no network here): answers a fixed status, keeps every request."""
from typing import Any, Dict, List


class MockTransport:
    """Records what would have gone on the wire."""

    def __init__(self, status: int = 200):
        self.status = status
        self.requests: List[Dict[str, Any]] = []

    def __call__(self, method, url, body, headers, timeout):
        self.requests.append({"method": method, "url": url, "body": body, "headers": dict(headers)})
        return self.status

    def bodies(self) -> List[bytes]:
        return [r["body"] for r in self.requests if r["method"] == "POST"]
