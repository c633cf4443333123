"""Worker for tests/test_multi_rank_cpu.py: one rank of a 2-process gloo job (no GPU).

Each rank generates its own uid-hash shard with the product's host generator, replays it (the C
oracle stands in for the engine, which needs a GPU), and the ranks exchange four counters the way
bench.py does.  Rank 0 then rebuilds the unsharded log and checks that the shards add up to it.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from k8s_watcher_b200 import generate_host, shard_of, sharding  # noqa: E402
from oracle import c_oracle  # noqa: E402

SEED, N_PER, PODS_PER = 77, 20_000, 900


def main():
    rank, world, local = sharding.rank_env()
    dist = sharding.init_process_group("gloo", local)
    assert (dist.get_rank(), dist.get_world_size()) == (rank, world)
    blob = sharding.broadcast_bytes(dist, bytes(range(128)) if rank == 0 else None)
    assert blob == bytes(range(128))

    cols = generate_host(SEED, N_PER, PODS_PER, dist=1, p_del=0.05, shard=rank, n_shards=world)
    assert all(shard_of(int(h), int(l), world) == rank for h, l in zip(cols["uid_hi"][:100], cols["uid_lo"][:100]))
    idx, st, dig = c_oracle.replay(cols, pods_hint=PODS_PER)
    counts = sharding.gather_counts_pg(dist, [st[k] for k in sharding.COUNT_NAMES])
    assert counts.shape == (world, 4) and int(counts[rank, 2]) == st["n_emitted"]
    slow = sharding.max_over_ranks(dist, [10.0 + rank, 5.0 - rank])
    assert slow == [10.0 + world - 1, 5.0]

    if rank == 0:
        # the union of all shards, merged back into rv order, replayed as ONE log
        parts = [generate_host(SEED, N_PER, PODS_PER, dist=1, p_del=0.05, shard=g, n_shards=world) for g in range(world)]
        allc = {k: np.concatenate([p[k] for p in parts]) for k in parts[0]}
        order = np.argsort(allc["rv"], kind="stable")
        allc = {k: v[order] for k, v in allc.items()}
        _, st_all, _ = c_oracle.replay(allc, pods_hint=PODS_PER * world)
        tot = counts.sum(axis=0)
        assert [int(x) for x in tot] == [st_all[k] for k in sharding.COUNT_NAMES], (tot, st_all)
        assert abs(sharding.throughput(int(tot[0]), 2.0) - int(tot[0]) / 2e-3) < 1e-6
        print("MULTI_RANK_OK", [int(x) for x in tot], flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
