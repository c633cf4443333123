"""Batches of up to 4 096 events take the one-launch path (csrc/pte_tiny.cuh); longer ones, and every batch under
PTE_TINY=0, the bulk kernels.  Both must give the oracle's answer -- one by one, mixed inside one stream, through
pte_stream_batch, and when a batch is refused.  (The reference's loop handles one event at a time,
watcher/pod_watcher.py:264-269; a stream cut into batches must notify exactly as the whole log replayed at once.)"""
import os

import numpy as np
import pytest

from k8s_watcher_b200 import PhaseTransitionEngine, PteError
from oracle import c_oracle

pytestmark = pytest.mark.gpu
SEED = 0x6B38735F77617463
FILTERS = {
    "staging": dict(critical_only=False, ns_allow=None),
    "development": dict(critical_only=False, ns_allow=[0, 1]),
    "production": dict(critical_only=True, ns_allow=[0, 1, 2, 3]),
}


@pytest.fixture(params=["one_launch", "bulk"])
def mode(request, monkeypatch):
    monkeypatch.setenv("PTE_TINY", "1" if request.param == "one_launch" else "0")
    return request.param


def _launches(mode, n):
    return 1 if (mode == "one_launch" and 0 < n <= 4096) else None


@pytest.mark.parametrize("env", list(FILTERS))
def test_small_logs_match_the_oracle(mode, env):
    flt = FILTERS[env]
    with PhaseTransitionEngine(5000, 5000) as e:
        e.set_filters(flt["ns_allow"], critical_only=flt["critical_only"])
        for n, pods, dist in [(1, 1, 0), (2, 1, 0), (31, 3, 1), (32, 32, 0), (33, 5, 1), (64, 2, 0), (65, 65, 0), (1000, 40, 1),
                              (2049, 2049, 0), (4095, 300, 1), (4096, 4096, 0), (4096, 7, 1), (4097, 100, 1)]:
            cols = c_oracle.generate(SEED + n, n, pods, dist=dist, p_del=0.1, p_nostatus=0.05, p_null=0.05)
            want, ost, dig = c_oracle.replay(cols, critical_only=flt["critical_only"], ns_allow=flt["ns_allow"], pods_hint=pods)
            e.ingest(cols)
            st = e.run(dedupe=True)
            got = e.fetch()
            assert np.array_equal(got.idx.astype(np.uint64), want), (n, pods)
            assert np.array_equal(got.type, cols["type"][want.astype(np.int64)]) and np.array_equal(got.rv, cols["rv"][want.astype(np.int64)])
            assert all(st[k] == ost[k] for k in ost), (n, st, ost)
            assert e.digest() == dig
            bits = np.unpackbits(e.fetch_bitmap().view(np.uint8), bitorder="little")[:n]
            assert np.array_equal(np.nonzero(bits)[0].astype(np.uint64), want)
            if _launches(mode, n):
                assert st["n_launches"] == 1
            elif n:
                assert st["n_launches"] > 1


@pytest.mark.parametrize("env", list(FILTERS))
def test_a_stream_of_small_and_large_batches_equals_one_replay(mode, env):
    """Batch sizes on both sides of the 4 096 limit inside ONE stream: the carried per-pod state has to survive the
    change of path in both directions."""
    n, pods = 60_000, 700
    flt = FILTERS[env]
    cols = c_oracle.generate(SEED + 9, n, pods, dist=1, p_del=0.08, p_nostatus=0.03, p_null=0.03)
    want, ost, _ = c_oracle.replay(cols, critical_only=flt["critical_only"], ns_allow=flt["ns_allow"], pods_hint=pods)
    rnd = np.random.default_rng(4)
    sizes = []
    while sum(sizes) < n:
        sizes.append(int(rnd.choice([1, 7, 64, 500, 4095, 4096, 4097, 9000, 3000])))
    edges = np.minimum(np.concatenate([[0], np.cumsum(sizes)]), n)
    got, tomb, kinds = [], 0, set()
    with PhaseTransitionEngine(10_000, pods) as e:
        e.set_filters(flt["ns_allow"], critical_only=flt["critical_only"])
        for k, (lo, hi) in enumerate(zip(edges[:-1], edges[1:])):
            if hi == lo:
                continue
            part = {c: v[lo:hi] for c, v in cols.items()}
            if k % 2:
                out, st = e.stream_batch(part)
                got.append(out.idx.astype(np.int64) + lo)
                assert np.array_equal(out.type, part["type"][out.idx.astype(np.int64)])
            else:
                e.ingest(part)
                st = e.run(dedupe=True, carry=True)
                got.append(e.fetch(columns=("idx",)).idx.astype(np.int64) + lo)
            tomb += st["n_tombstones"]
            kinds.add(st["n_launches"] == 1)
        assert np.array_equal(np.concatenate(got), want.astype(np.int64))
        assert tomb == ost["n_tombstones"]
        assert kinds == ({True, False} if mode == "one_launch" else {False})


def test_refused_small_batches_leave_the_stream_as_it_was(mode):
    n, pods = 9000, 200
    cols = c_oracle.generate(SEED + 2, n, pods, dist=0, p_del=0.05)
    want, _, _ = c_oracle.replay(cols, pods_hint=pods)
    cut = [0, 3000, 6000, 9000]
    with PhaseTransitionEngine(4096, pods) as e:
        out0, _ = e.stream_batch({c: v[cut[0]:cut[1]] for c, v in cols.items()})
        # 1. resourceVersion goes backwards inside the batch
        bad = {c: v[cut[1]:cut[2]].copy() for c, v in cols.items()}
        bad["rv"][10] = bad["rv"][9] - 1
        with pytest.raises(PteError) as ei:
            e.stream_batch(bad)
        assert ei.value.code == -7
    # (an unordered batch ends the stream: start again and test the refusals that do not)
    with PhaseTransitionEngine(4096, pods) as e:
        got = []
        out, _ = e.stream_batch({c: v[cut[0]:cut[1]] for c, v in cols.items()})
        got.append(out.idx.astype(np.int64))
        # 2. an event code out of range, 3. the all-zero uid: refused, and the columns of the refused batch can be read back
        for column, value in (("phase", 77), ("uid", 0)):
            bad = {c: v[cut[1]:cut[2]].copy() for c, v in cols.items()}
            if column == "uid":
                bad["uid_hi"][5] = 0
                bad["uid_lo"][5] = 0
            else:
                bad[column][5] = value
            with pytest.raises(PteError):
                e.stream_batch(bad)
            back = e.read_log()
            assert all(np.array_equal(back[c], bad[c]) for c in bad)
    # a refused batch invalidates the stream (as in the bulk path); what must hold is that a NEW stream is clean
    with PhaseTransitionEngine(4096, pods) as e:
        got = []
        for lo, hi in zip(cut[:-1], cut[1:]):
            out, _ = e.stream_batch({c: v[lo:hi] for c, v in cols.items()})
            got.append(out.idx.astype(np.int64) + lo)
        assert np.array_equal(np.concatenate(got), want.astype(np.int64))


def test_dictionary_rebuild_inside_a_stream_of_small_batches(mode):
    """Far more distinct uids than max_pods over the life of the stream, few alive at a time: the dictionary is rebuilt
    from the live pods when it fills up (proactively and after a refused batch) and the stream goes on."""
    max_pods, batch, rounds = 512, 300, 60
    rnd = np.random.default_rng(8)
    rows = []
    rv = 1
    alive = []
    next_pod = 1
    for _ in range(rounds * batch):
        r = rnd.random()
        if alive and (r < 0.45 or len(alive) > max_pods // 3):
            p = alive.pop(int(rnd.integers(len(alive))))
            rows.append((p, rv, 2, int(rnd.integers(5))))
        elif alive and r < 0.75:
            p = alive[int(rnd.integers(len(alive)))]
            rows.append((p, rv, 1, int(rnd.integers(5))))
        else:
            p = next_pod
            next_pod += 1
            alive.append(p)
            rows.append((p, rv, 0, int(rnd.integers(5))))
        rv += 1
    a = np.array(rows, dtype=np.uint64)
    cols = {"uid_hi": a[:, 0] * np.uint64(0x9E3779B97F4A7C15), "uid_lo": a[:, 0] ^ np.uint64(0x5555), "rv": a[:, 1],
            "ns_id": np.zeros(len(a), np.uint32), "phase": a[:, 3].astype(np.uint8), "type": a[:, 2].astype(np.uint8)}
    assert next_pod > 10 * max_pods
    want, _, _ = c_oracle.replay(cols)
    got = []
    with PhaseTransitionEngine(batch, max_pods) as e:
        for lo in range(0, len(a), batch):
            out, st = e.stream_batch({c: v[lo:lo + batch] for c, v in cols.items()})
            got.append(out.idx.astype(np.int64) + lo)
        assert st["n_rebuilds"] > 0
    assert np.array_equal(np.concatenate(got), want.astype(np.int64))


@pytest.mark.parametrize("env", list(FILTERS))
def test_small_unordered_log_replays_in_rv_order(mode, env):
    """A short log that is not in resourceVersion order is pre-sorted on the device whichever path replays it; the
    engine's copy of the log stays as it was handed over."""
    n, pods = 3000, 120
    flt = FILTERS[env]
    cols = c_oracle.generate(SEED + 11, n, pods, dist=1, p_del=0.07, p_nostatus=0.03, p_null=0.03)
    perm = np.random.default_rng(6).permutation(n)
    shuf = {k: v[perm] for k, v in cols.items()}
    want, ost, _ = c_oracle.replay(cols, critical_only=flt["critical_only"], ns_allow=flt["ns_allow"], pods_hint=pods)
    with PhaseTransitionEngine(4096, pods) as e:
        e.set_filters(flt["ns_allow"], critical_only=flt["critical_only"])
        e.ingest(shuf)
        st = e.run(dedupe=True)
        assert st["ordered"] == 0 and st["n_emitted"] == ost["n_emitted"]
        got = e.fetch(columns=("idx",)).idx.astype(np.int64)
        assert np.array_equal(perm[got], want.astype(np.int64))
        back = e.read_log()
        assert all(np.array_equal(back[k], shuf[k]) for k in shuf)


@pytest.mark.parametrize("tail", [1, 1000, 4096])
def test_pipelined_replay_whose_last_chunk_is_small(mode, tail):
    """pte_replay_pipelined cuts the log into carried chunks; a short last chunk takes the one-launch path."""
    n, pods = 3 * 8192 + tail, 400
    cols = c_oracle.generate(SEED + 13, n, pods, dist=1, p_del=0.06)
    want, _, _ = c_oracle.replay(cols, pods_hint=pods)
    with PhaseTransitionEngine(n, pods) as e:
        for with_rv in (True, False):
            words, st = e.replay_pipelined(cols, chunk_events=8192, with_rv=with_rv)
            bits = np.unpackbits(words.view(np.uint8), bitorder="little")[:n]
            assert np.array_equal(np.nonzero(bits)[0].astype(np.uint64), want), with_rv
            assert st["n_emitted"] == len(want)
