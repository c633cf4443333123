"""Integer codes of the replay-log format (mirror of include/pte.h)."""
from __future__ import annotations

TYPE_NAMES = ("ADDED", "MODIFIED", "DELETED")
TYPE_ADDED, TYPE_MODIFIED, TYPE_DELETED = 0, 1, 2
TYPE_CODE = {n: i for i, n in enumerate(TYPE_NAMES)}

PHASE_NAMES = ("Pending", "Running", "Succeeded", "Failed", "Unknown")
PH_PENDING, PH_RUNNING, PH_SUCCEEDED, PH_FAILED, PH_UNKNOWN = range(5)
PH_NULL = 5            # status present, status.phase is None
PH_NOSTATUS = 6        # pod.status is None -> payload says "Unknown" (reference pod_watcher.py:167)
PH_FIRST_DYNAMIC = 7   # other phase strings are interned per log from here
PH_MAX = 62

DEFAULT_NAMESPACES = ("default", "kube-system", "production", "monitoring", "other")


def uid_to_str(hi: int, lo: int) -> str:
    """128-bit uid -> the 8-4-4-4-12 hex form of metadata.uid."""
    h = "%016x%016x" % (int(hi), int(lo))
    return "-".join((h[0:8], h[8:12], h[12:16], h[16:20], h[20:32]))


def str_to_uid(uid: str):
    """metadata.uid -> (hi, lo).  UUID-shaped strings map bijectively; anything else is
    hashed to 128 bits (blake2b) -- callers that need exactness for such uids keep a dictionary."""
    s = uid.replace("-", "")
    if len(s) == 32:
        try:
            v = int(s, 16)
            return v >> 64, v & 0xFFFFFFFFFFFFFFFF
        except ValueError:
            pass
    import hashlib
    v = int.from_bytes(hashlib.blake2b(uid.encode(), digest_size=16).digest(), "big")
    return v >> 64, v & 0xFFFFFFFFFFFFFFFF
