"""ctypes binding of include/pte.h.  Loads k8s_watcher_b200/libpte.so and nothing else:
there is no Python or CPU implementation behind these calls."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpte.so")

PTE_OK = 0
STATUS_NAMES = {
    0: "PTE_OK", -1: "PTE_ERR_INVALID", -2: "PTE_ERR_CUDA", -3: "PTE_ERR_NO_DEVICE", -4: "PTE_ERR_CAPACITY",
    -5: "PTE_ERR_TABLE_FULL", -6: "PTE_ERR_BAD_EVENT", -7: "PTE_ERR_UNORDERED", -8: "PTE_ERR_STATE",
    -9: "PTE_ERR_NCCL", -10: "PTE_ERR_PARSE",
}


class PteError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__("%s (%d): %s" % (STATUS_NAMES.get(code, "PTE_ERR"), code, message))
        self.code = code


class PteStats(C.Structure):
    _fields_ = [
        ("n_in", C.c_uint64), ("n_survived", C.c_uint64), ("n_emitted", C.c_uint64), ("n_tombstones", C.c_uint64),
        ("n_pods", C.c_uint64), ("ordered", C.c_uint32), ("n_launches", C.c_uint32),
        ("ms_total", C.c_float), ("ms_intern", C.c_float), ("ms_sort_key", C.c_float), ("ms_diff", C.c_float),
        ("ms_sort_order", C.c_float), ("ms_assemble", C.c_float), ("ms_pass_avg", C.c_float),
        ("n_passes", C.c_uint32), ("n_diff_rounds", C.c_uint32), ("path", C.c_uint32), ("n_pending", C.c_uint32),
        ("ms_part_hist", C.c_float), ("ms_part_scatter", C.c_float), ("ms_part_replay", C.c_float),
        ("n_rebuilds", C.c_uint32), ("ms_rebuild", C.c_float),
    ]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class PteGenParams(C.Structure):
    _fields_ = [
        ("seed", C.c_uint64), ("n_events", C.c_uint64), ("n_pods", C.c_uint64), ("dist", C.c_uint32),
        ("p_del_q24", C.c_uint32), ("p_nostatus_q24", C.c_uint32), ("p_null_q24", C.c_uint32),
        ("shard", C.c_uint32), ("n_shards", C.c_uint32),
    ]


_u64p, _u32p, _u8p = C.POINTER(C.c_uint64), C.POINTER(C.c_uint32), C.POINTER(C.c_uint8)

# name -> (restype, argtypes): every symbol include/pte.h declares
SYMBOLS = {
    "pte_device_count": (C.c_int, []),
    "pte_create": (C.c_int, [C.c_int, C.c_uint64, C.c_uint64, C.POINTER(C.c_void_p)]),
    "pte_destroy": (None, [C.c_void_p]),
    "pte_last_error": (C.c_char_p, [C.c_void_p]),
    "pte_set_filters": (C.c_int, [C.c_void_p, _u32p, C.c_int, C.c_int, C.c_int]),
    "pte_ingest": (C.c_int, [C.c_void_p, _u64p, _u64p, _u64p, _u32p, _u8p, _u8p, C.c_uint64]),
    "pte_ingest_json": (C.c_int, [C.c_void_p, C.c_char_p, C.c_uint64, _u64p, _u64p]),
    "pte_payload_timing": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), _u64p]),
    "pte_json_timing": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float), _u64p]),
    "pte_json_stream": (C.c_int, [C.c_void_p, C.c_int]),
    "pte_json_strings": (C.c_int, [C.c_void_p, C.c_int, C.c_char_p, C.c_uint64, _u32p, C.c_uint32, _u32p]),
    "pte_generate": (C.c_int, [C.c_void_p, C.POINTER(PteGenParams)]),
    "pte_generate_host": (C.c_int, [C.POINTER(PteGenParams), _u64p, _u64p, _u64p, _u32p, _u8p, _u8p]),
    "pte_read_log": (C.c_int, [C.c_void_p, _u64p, _u64p, _u64p, _u32p, _u8p, _u8p, C.c_uint64, _u64p]),
    "pte_run": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(PteStats)]),
    "pte_fetch": (C.c_int, [C.c_void_p, _u32p, _u64p, _u64p, _u32p, _u8p, _u8p, C.c_uint64, _u64p]),
    "pte_fetch_bitmap": (C.c_int, [C.c_void_p, _u32p, C.c_uint64, _u64p]),
    "pte_grow": (C.c_int, [C.c_void_p, C.c_uint64]),
    "pte_stream_batch": (C.c_int, [C.c_void_p, _u64p, _u64p, _u64p, _u32p, _u8p, _u8p, C.c_uint64, C.c_int, _u32p, _u8p,
                                   C.c_uint64, C.POINTER(PteStats)]),
    "pte_replay_pipelined": (C.c_int, [C.c_void_p, _u64p, _u64p, _u64p, _u32p, _u8p, _u8p, C.c_uint64, C.c_uint64, C.c_int,
                                       _u32p, C.POINTER(PteStats)]),
    "pte_digest": (C.c_int, [C.c_void_p, _u64p]),
    "pte_serialize_payloads": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint64, C.c_char_p, C.c_uint32, _u32p, C.c_uint32,
                                         C.c_uint32, C.c_void_p, C.c_uint64, _u64p, _u64p]),
    "pte_host_alloc": (C.c_void_p, [C.c_size_t]),
    "pte_host_free": (None, [C.c_void_p]),
    "pte_shard_of": (C.c_uint32, [C.c_uint64, C.c_uint64, C.c_uint32]),
    "pte_debug_sort_u64": (C.c_int, [C.c_void_p, _u64p, C.c_uint64, C.c_int, C.c_int]),
    "pte_partition_u64_static": (C.c_int, [_u64p, _u64p, C.c_uint64, C.c_int, C.c_int, C.c_void_p, C.c_char_p, C.c_size_t]),
    "pte_comm_unique_id": (C.c_int, [C.c_char_p]),
    "pte_comm_init": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int, C.c_int]),
    "pte_gather_counts": (C.c_int, [C.c_void_p, _u64p]),
}

_LIB = None


def lib() -> C.CDLL:
    """The loaded library.  Raises if it has not been built (python __graft_entry__.py build)."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                "k8s_watcher_b200: %s is missing -- build it with `make -C k8s_watcher_b200/csrc` "
                "(there is no fallback implementation)" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)       # AttributeError if the .so is stale
            fn.restype, fn.argtypes = res, args
        _LIB = L
    return _LIB


def device_count() -> int:
    return int(lib().pte_device_count())


def last_error(handle) -> str:
    msg = lib().pte_last_error(handle)
    return msg.decode(errors="replace") if msg else ""


def check(rc: int, handle=None) -> None:
    if rc != PTE_OK:
        raise PteError(rc, last_error(handle))
