"""ctypes front for oracle/liboracle.so (see oracle.c).  TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Dict, Iterable, Optional, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build() -> str:
    subprocess.run(["make", "-s", "-C", _HERE], check=True)
    return os.path.join(_HERE, "liboracle.so")


def lib() -> C.CDLL:
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        u64p, u32p, u8p = C.POINTER(C.c_uint64), C.POINTER(C.c_uint32), C.POINTER(C.c_uint8)
        L.orc_generate.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32,
                                   C.c_uint32, C.c_uint32, C.c_uint32, u64p, u64p, u64p, u32p, u8p, u8p]
        L.orc_generate.restype = None
        L.orc_replay.argtypes = [u64p, u64p, u64p, u32p, u8p, u8p, C.c_uint64, C.c_int, u32p, C.c_int, C.c_int,
                                 C.c_uint64, u64p, u64p]
        L.orc_replay.restype = C.c_uint64
        L.orc_generate_replay.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32,
                                          C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, u32p, C.c_int, C.c_int, u64p]
        L.orc_generate_replay.restype = C.c_uint64
        L.orc_unmix64.argtypes = [C.c_uint64]
        L.orc_unmix64.restype = C.c_uint64
        L.orc_shard_of.argtypes = [C.c_uint64, C.c_uint32]
        L.orc_shard_of.restype = C.c_uint32
        _LIB = L
    return _LIB


def _p(a: np.ndarray, t):
    return a.ctypes.data_as(C.POINTER(t))


def q24(p: float) -> int:
    return int(round(p * (1 << 24)))


def generate(seed: int, n: int, n_pods_local: int, *, dist: int = 0, p_del: float = 0.0, p_nostatus: float = 0.0,
             p_null: float = 0.0, shard: int = 0, n_shards: int = 1) -> Dict[str, np.ndarray]:
    cols = {
        "uid_hi": np.empty(n, np.uint64), "uid_lo": np.empty(n, np.uint64), "rv": np.empty(n, np.uint64),
        "ns_id": np.empty(n, np.uint32), "phase": np.empty(n, np.uint8), "type": np.empty(n, np.uint8),
    }
    lib().orc_generate(seed, n, n_pods_local, dist, q24(p_del), q24(p_nostatus), q24(p_null), shard, n_shards,
                       _p(cols["uid_hi"], C.c_uint64), _p(cols["uid_lo"], C.c_uint64), _p(cols["rv"], C.c_uint64),
                       _p(cols["ns_id"], C.c_uint32), _p(cols["phase"], C.c_uint8), _p(cols["type"], C.c_uint8))
    return cols


def _allow(ns_allow: Optional[Iterable[int]]):
    a = np.ascontiguousarray(sorted(int(x) for x in ns_allow) if ns_allow else [], dtype=np.uint32)
    return a, (_p(a, C.c_uint32) if a.size else None), int(a.size)


def replay(cols: Dict[str, np.ndarray], *, critical_only: bool = False, ns_allow: Optional[Iterable[int]] = None,
           dedupe: bool = True, pods_hint: int = 0) -> Tuple[np.ndarray, Dict[str, int], int]:
    """-> (emitted arrival indices u64, stats, digest)."""
    n = int(cols["rv"].shape[0])
    c = {k: np.ascontiguousarray(v) for k, v in cols.items()}
    out = np.empty(max(n, 1), np.uint64)
    stats = np.zeros(4, np.uint64)
    a, ap, na = _allow(ns_allow)
    dig = lib().orc_replay(_p(c["uid_hi"], C.c_uint64), _p(c["uid_lo"], C.c_uint64), _p(c["rv"], C.c_uint64),
                           _p(c["ns_id"], C.c_uint32), _p(c["phase"], C.c_uint8), _p(c["type"], C.c_uint8),
                           n, int(critical_only), ap, na, int(dedupe), pods_hint, _p(out, C.c_uint64),
                           _p(stats, C.c_uint64))
    st = dict(zip(("n_in", "n_survived", "n_emitted", "n_tombstones"), (int(x) for x in stats)))
    return out[: st["n_emitted"]].copy(), st, int(dig)


def generate_replay(seed: int, n: int, n_pods_local: int, *, dist: int = 0, p_del: float = 0.0,
                    p_nostatus: float = 0.0, p_null: float = 0.0, shard: int = 0, n_shards: int = 1,
                    critical_only: bool = False, ns_allow: Optional[Iterable[int]] = None,
                    dedupe: bool = True) -> Tuple[Dict[str, int], int]:
    stats = np.zeros(4, np.uint64)
    a, ap, na = _allow(ns_allow)
    dig = lib().orc_generate_replay(seed, n, n_pods_local, dist, q24(p_del), q24(p_nostatus), q24(p_null), shard,
                                    n_shards, int(critical_only), ap, na, int(dedupe), _p(stats, C.c_uint64))
    st = dict(zip(("n_in", "n_survived", "n_emitted", "n_tombstones"), (int(x) for x in stats)))
    return st, int(dig)
