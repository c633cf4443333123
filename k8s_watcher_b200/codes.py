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


_CANONICAL_UID = None


def str_to_uid(uid: str):
    """metadata.uid -> (hi, lo) for UUID-shaped strings (32 hex digits, dashes ignored: bijective on the canonical
    lower-case 8-4-4-4-12 form Kubernetes emits).  Anything else raises ValueError: such uids need a UidInterner."""
    s = uid.replace("-", "")
    if len(s) == 32:
        try:
            v = int(s, 16)
            return v >> 64, v & 0xFFFFFFFFFFFFFFFF
        except ValueError:
            pass
    raise ValueError("metadata.uid %r is not a 128-bit hex uid" % (uid,))


class UidInterner:
    """EXACT metadata.uid -> 128-bit engine key for whatever strings arrive (the reference takes any string,
    watcher/pod_watcher.py:164).

    The canonical form Kubernetes emits (lower-case hex, 8-4-4-4-12) maps to its own 128-bit value.  Every other
    string -- and the few canonical ones whose value falls into the reserved block ``hi == 2**64 - 1`` or is zero -- gets
    the next number of the reserved block from a dictionary, so two different strings never share a key and one string
    always gets the same key for as long as the interner lives (one per watch stream)."""

    RESERVED_HI = 0xFFFFFFFFFFFFFFFF

    def __init__(self):
        import re
        self._canonical = re.compile(r"[0-9a-f]{8}-[0-9a-f]{4}-[0-9a-f]{4}-[0-9a-f]{4}-[0-9a-f]{12}\Z")
        self._ids = {}          # string -> lo of its reserved key
        self._strings = []      # lo - 1 -> string

    def key(self, uid: str):
        if self._canonical.match(uid):
            v = int(uid.replace("-", ""), 16)
            hi, lo = v >> 64, v & 0xFFFFFFFFFFFFFFFF
            if hi != self.RESERVED_HI and v != 0:
                return hi, lo
        lo = self._ids.get(uid)
        if lo is None:
            self._strings.append(uid)
            lo = self._ids[uid] = len(self._strings)
        return self.RESERVED_HI, lo

    def string(self, hi: int, lo: int) -> str:
        """The uid string of a key this interner handed out (or the canonical form of a direct one)."""
        if int(hi) == self.RESERVED_HI and 0 < int(lo) <= len(self._strings):
            return self._strings[int(lo) - 1]
        return uid_to_str(hi, lo)

    def __len__(self):
        return len(self._strings)
