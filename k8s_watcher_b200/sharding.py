"""Host-side plumbing for one-engine-per-GPU runs (one process per GPU, torchrun-style environment).

The data path never crosses GPUs: a log is split by uid hash (``pte_shard_of``), every rank replays
its own shard, and the only exchange is four counters per rank.  ``torch.distributed`` is used for
the rendezvous only (it is what launches the ranks); on GPUs the counters themselves travel through
``pte_gather_counts`` (NCCL inside libpte.so), on CPU-only hosts (tests) through the process group.
"""
from __future__ import annotations

import os
from typing import List, Optional, Sequence, Tuple

import numpy as np

COUNT_NAMES = ("n_in", "n_survived", "n_emitted", "n_tombstones")


def rank_env() -> Tuple[int, int, int]:
    """(rank, world_size, local_rank) from the launcher's environment; (0, 1, 0) when not launched."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def init_process_group(backend: str, local_rank: int):
    """Join the launcher's rendezvous (MASTER_ADDR/MASTER_PORT from the environment)."""
    import torch
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if backend == "nccl":
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    else:
        dist.init_process_group(backend)
    return dist


def broadcast_bytes(dist, payload: Optional[bytes], src: int = 0) -> bytes:
    """Ship a small blob (the 128-byte NCCL unique id) from ``src`` to every rank."""
    box = [payload if dist.get_rank() == src else None]
    dist.broadcast_object_list(box, src=src)
    return box[0]


def gather_counts_pg(dist, mine: Sequence[int]) -> np.ndarray:
    """All-gather of the four per-rank counters over the process group -> array [world, 4]."""
    import torch
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.tensor([int(x) for x in mine], dtype=torch.int64, device=dev)
    out = [torch.zeros_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return np.stack([o.cpu().numpy() for o in out]).astype(np.uint64)


def max_over_ranks(dist, values: Sequence[float]) -> List[float]:
    """Element-wise maximum over ranks (timings are always reported as the slowest rank's)."""
    import torch
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.tensor(list(values), dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(x) for x in t.cpu()]


def throughput(total_events: int, step_ms_max: float) -> float:
    """Whole-job events/s: all ranks' events over the slowest rank's step time."""
    return total_events / (step_ms_max * 1e-3)


def bind_to_gpu_numa(local_rank: int) -> Optional[dict]:
    """Pin this process to the CPUs of the NUMA node its GPU hangs off (so that pinned staging buffers, allocated
    afterwards, are first-touched there and host -> device copies do not cross sockets).  Best effort: returns what
    it did, or None when the topology cannot be read."""
    import subprocess
    try:
        out = subprocess.run(["nvidia-smi", "-i", str(local_rank), "--query-gpu=pci.bus_id", "--format=csv,noheader"],
                             capture_output=True, text=True, timeout=10).stdout.strip()
        bus = out.lower()
        if bus.startswith("00000000:"):
            bus = "0000:" + bus.split(":", 1)[1]
        with open("/sys/bus/pci/devices/%s/numa_node" % bus) as f:
            node = int(f.read().strip())
        if node < 0:
            return {"gpu_bus": bus, "numa_node": node, "bound": False}
        with open("/sys/devices/system/node/node%d/cpulist" % node) as f:
            cpulist = f.read().strip()
        cpus = set()
        for part in cpulist.split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = cpus & set(os.sched_getaffinity(0))
        if allowed:
            os.sched_setaffinity(0, allowed)
        return {"gpu_bus": bus, "numa_node": node, "cpus": cpulist, "bound": bool(allowed)}
    except Exception:
        return None
