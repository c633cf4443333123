"""B200 bulk pod phase-transition engine behind the k8s-watcher ``PodWatcher`` surface.

Python here is the host shim only; the path itself is CUDA (csrc/ -> libpte.so) reached
through the C ABI in include/pte.h.  Importing the package does not need a GPU; creating an
engine does, and fails loudly without one.
"""
from .clusterapi_client import ClusterApiClient  # noqa: F401
from .engine import EmittedColumns, PhaseTransitionEngine, PteError, generate_host, shard_of  # noqa: F401
from .pod_watcher import Batch, PodWatcher  # noqa: F401
from .replay_log import ReplayLog  # noqa: F401
from . import sharding  # noqa: F401

__all__ = ["PodWatcher", "Batch", "ClusterApiClient", "PhaseTransitionEngine", "EmittedColumns",
           "ReplayLog", "PteError", "generate_host", "shard_of"]
