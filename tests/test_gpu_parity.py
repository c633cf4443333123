"""Parity of the CUDA path (through the C ABI) with the oracle.  Needs a B200: -m gpu."""
import glob
import gzip
import json
import os

import numpy as np
import pytest

import fake_kubernetes as fk
from mock_transport import MockTransport
from k8s_watcher_b200 import PhaseTransitionEngine, PodWatcher, PteError, ReplayLog, generate_host
from oracle import c_oracle
from oracle import podwatch_oracle as po

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "ref_*.json.gz")))
SEED = 0x6B38735F77617463


@pytest.fixture(scope="module")
def eng():
    e = PhaseTransitionEngine(4_000_000, 300_000)
    yield e
    e.close()


# ---- the sort in isolation ---------------------------------------------------
@pytest.mark.parametrize("n", [0, 1, 2, 31, 32, 33, 255, 4095, 4096, 4097, 8191, 65_537, 1_000_003, 3_999_999])
def test_radix_sort_matches_numpy_stable(eng, n):
    rng = np.random.default_rng(n + 1)
    rec = rng.integers(0, 1 << 63, n, dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, n, dtype=np.uint64)
    for lo, hi in ((0, 64), (8, 38), (38, 62), (5, 6), (17, 17), (40, 47)):
        got = eng.debug_sort(rec, lo, hi)
        field = (rec >> np.uint64(lo)) & np.uint64((1 << (hi - lo)) - 1) if hi > lo else np.zeros(n, np.uint64)
        want = rec[np.argsort(field, kind="stable")]
        assert np.array_equal(got, want), (n, lo, hi)


@pytest.mark.parametrize("kind", ["constant", "two_values", "sorted", "reversed", "hot_bin"])
def test_radix_sort_skewed_inputs(eng, kind):
    n = 700_001
    rng = np.random.default_rng(5)
    low = rng.integers(0, 1 << 20, n, dtype=np.uint64)      # payload bits that must keep their order
    if kind == "constant":
        key = np.full(n, 77, np.uint64)
    elif kind == "two_values":
        key = rng.integers(0, 2, n, dtype=np.uint64) * np.uint64(0xABCDEF)
    elif kind == "sorted":
        key = np.arange(n, dtype=np.uint64)
    elif kind == "reversed":
        key = np.arange(n, dtype=np.uint64)[::-1].copy()
    else:
        key = np.where(rng.random(n) < 0.9, np.uint64(0x1234), rng.integers(0, 1 << 24, n, dtype=np.uint64))
    rec = (key << np.uint64(20)) | low
    got = eng.debug_sort(rec, 20, 44)
    want = rec[np.argsort(key, kind="stable")]
    assert np.array_equal(got, want)


# ---- the generator -----------------------------------------------------------
@pytest.mark.parametrize("kw", [
    dict(n=1000, pods=50, dist=0, p_del=0.05),
    dict(n=300_000, pods=7_000, dist=0, p_del=0.0),
    dict(n=300_000, pods=7_000, dist=1, p_del=0.05, p_nostatus=0.02, p_null=0.02),
    dict(n=100_000, pods=1, dist=1, p_del=0.3),
    dict(n=50_000, pods=9_000, dist=0, p_del=0.05, shard=3, n_shards=8),
])
def test_device_generator_matches_host_and_oracle(eng, kw):
    kw = dict(kw)
    n, pods = kw.pop("n"), kw.pop("pods")
    eng.generate(SEED, n, pods, **kw)
    dev = eng.read_log()
    host = generate_host(SEED, n, pods, **kw)
    orc = c_oracle.generate(SEED, n, pods, **kw)
    for k in orc:
        assert np.array_equal(host[k], orc[k]), ("host", k)
        assert np.array_equal(dev[k], orc[k]), ("device", k)


# ---- replay parity ------------------------------------------------------------
FILTERS = {
    "staging": dict(critical_only=False, ns_allow=None),
    "development": dict(critical_only=False, ns_allow=[0, 1]),
    "production": dict(critical_only=True, ns_allow=[0, 1, 2, 3]),
}


def run_engine(eng, cols, flt, dedupe):
    eng.set_filters(flt["ns_allow"], critical_only=flt["critical_only"])
    eng.ingest(cols)
    st = eng.run(dedupe=dedupe)
    return eng.fetch(), st


def check_against_oracle(eng, cols, flt, dedupe, pods):
    got, st = run_engine(eng, cols, flt, dedupe)
    idx, ost, dig = c_oracle.replay(cols, critical_only=flt["critical_only"], ns_allow=flt["ns_allow"], dedupe=dedupe,
                                    pods_hint=pods)
    assert {k: st[k] for k in ost} == ost
    assert np.array_equal(got.idx.astype(np.uint64), idx)
    i = idx.astype(np.int64)
    assert np.array_equal(got.rv, cols["rv"][i])
    assert np.array_equal(got.ns_id, cols["ns_id"][i])
    assert np.array_equal(got.phase, cols["phase"][i])
    assert np.array_equal(got.type, cols["type"][i])
    # the interned key is a bijection of the uid (keys exist only when the run interned, i.e. dedupe)
    if len(i) and dedupe:
        uid = cols["uid_hi"][i].astype(object) * (1 << 64) + cols["uid_lo"][i].astype(object)
        pairs = set(zip(uid.tolist(), got.key.tolist()))
        assert len(pairs) == len(set(uid.tolist())) == len(set(got.key.tolist()))
    assert eng.digest() == dig
    return st


@pytest.mark.parametrize("env", list(FILTERS))
@pytest.mark.parametrize("dedupe", [True, False])
@pytest.mark.parametrize("shape", [
    dict(n=1000, pods=50, dist=0, p_del=0.05),                              # BASELINE config 1
    dict(n=250_000, pods=4_000, dist=0, p_del=0.0),
    dict(n=250_000, pods=4_000, dist=1, p_del=0.05, p_nostatus=0.03, p_null=0.03),
])
def test_replay_matches_oracle(eng, env, dedupe, shape):
    shape = dict(shape)
    n, pods = shape.pop("n"), shape.pop("pods")
    cols = c_oracle.generate(SEED + 1, n, pods, **shape)
    check_against_oracle(eng, cols, FILTERS[env], dedupe, pods)


@pytest.mark.parametrize("n,pods", [(1, 1), (2, 1), (3, 2), (1023, 5), (1024, 5), (1025, 5), (4097, 4097)])
def test_ragged_sizes(eng, n, pods):
    cols = c_oracle.generate(n * 31 + pods, n, pods, dist=n % 2, p_del=0.2, p_nostatus=0.1, p_null=0.1)
    for env in FILTERS:
        check_against_oracle(eng, cols, FILTERS[env], True, pods)


def test_empty_log(eng):
    cols = c_oracle.generate(1, 0, 10)
    got, st = run_engine(eng, cols, FILTERS["staging"], True)
    assert len(got) == 0 and st["n_emitted"] == 0 and eng.digest() == 0


def test_everything_filtered_out(eng):
    cols = c_oracle.generate(2, 5000, 100)
    eng.set_filters([], ns_filter_on=True)          # allow-list names that never occur in the log
    eng.ingest(cols)
    st = eng.run()
    assert st["n_survived"] == 0 and st["n_emitted"] == 0 and len(eng.fetch()) == 0


def test_single_pod_all_same_phase(eng):
    n = 100_000
    cols = c_oracle.generate(3, n, 1)
    cols["phase"][:] = po.PH_RUNNING
    got, st = run_engine(eng, cols, FILTERS["staging"], True)
    assert st["n_emitted"] == 1 and got.idx.tolist() == [0]


def test_all_deleted_are_tombstones(eng):
    n = 50_000
    cols = c_oracle.generate(4, n, 300)
    cols["type"][:] = po.TYPE_DELETED
    got, st = run_engine(eng, cols, FILTERS["production"], True)
    keep = np.isin(cols["ns_id"], [0, 1, 2, 3])
    assert st["n_emitted"] == st["n_tombstones"] == int(keep.sum())
    assert np.array_equal(got.idx, np.nonzero(keep)[0].astype(np.uint32))


def test_statusless_and_unknown_share_an_effective_phase(eng):
    # NOSTATUS then Unknown for the same pod: the payload says "Unknown" twice -> one notification
    cols = c_oracle.generate(5, 4, 1)
    cols["type"][:] = [0, 1, 1, 1]
    cols["phase"][:] = [po.PH_NOSTATUS, po.PH_UNKNOWN, po.PH_NOSTATUS, po.PH_RUNNING]
    got, st = run_engine(eng, cols, FILTERS["staging"], True)
    assert got.idx.tolist() == [0, 3]


def test_dictionary_is_exact_for_adversarial_uids(eng):
    # uids that differ in one half only, and uids with a zero half
    n = 60_000
    rng = np.random.default_rng(11)
    base = rng.integers(1, 1 << 62, 200, dtype=np.uint64)
    pick = rng.integers(0, 200, n)
    cols = c_oracle.generate(6, n, 200)
    cols["uid_hi"][:] = np.where(pick % 3 == 0, np.uint64(0), base[pick])
    cols["uid_lo"][:] = np.where(pick % 3 == 1, np.uint64(0), base[(pick * 7) % 200])
    cols["uid_lo"][(cols["uid_hi"] == 0) & (cols["uid_lo"] == 0)] = 9
    check_against_oracle(eng, cols, FILTERS["staging"], True, 1000)


def test_config2_full_parity():
    """BASELINE config 2: 10 M events / 100 k pods, uniform churn -- every emitted record compared."""
    n, pods = 10_000_000, 100_000
    with PhaseTransitionEngine(n, pods) as e:
        e.generate(SEED, n, pods)
        st = e.run()
        got = e.fetch(columns=("idx", "type", "phase"))
        cols = c_oracle.generate(SEED, n, pods)
        idx, ost, dig = c_oracle.replay(cols, pods_hint=pods)
        assert np.array_equal(got.idx.astype(np.uint64), idx)
        assert {k: st[k] for k in ost} == ost and e.digest() == dig
        assert 0.79 < st["n_emitted"] / n < 0.82          # f = h + (1-h) * 0.8


def test_zipf_with_tombstones_digest_100m():
    """Shape of BASELINE config 4 at 100 M events: octave-Zipf pod activity, 5 % delete draws.
    Too big to compare record by record here, so counts + order-sensitive digest."""
    n, pods = 100_000_000, 1_000_000
    with PhaseTransitionEngine(n, pods) as e:
        e.generate(SEED, n, pods, dist=1, p_del=0.05)
        st = e.run()
        ost, dig = c_oracle.generate_replay(SEED, n, pods, dist=1, p_del=0.05)
        assert {k: st[k] for k in ost} == ost
        assert e.digest() == dig
        # idempotence: the same log replays to the same answer
        st2 = e.run()
        assert st2["n_emitted"] == st["n_emitted"] and e.digest() == dig


# ---- the diff stage's bitmap slices (1 B events need two; PTE_DIFF_ROUND_MB shrinks them for the test) -------
@pytest.mark.parametrize("mode", ["deferred", "looped", "four_slices"])
@pytest.mark.parametrize("shape", [
    dict(dist=0, p_del=0.0),                         # 80 % emitted: the diff clears the suppressed bits
    dict(dist=1, p_del=0.05, quiet=True),            # almost nothing emitted: the diff sets the emitted bits
])
def test_bitmap_slices_give_the_same_answer(monkeypatch, mode, shape):
    monkeypatch.setenv("PTE_DIFF_ROUND_MB", "1")                       # 8 M positions per slice
    if mode == "looped":
        monkeypatch.setenv("PTE_NO_DEFER", "1")
    n, pods = (30_000_000 if mode == "four_slices" else 12_000_000), 150_000
    shape = dict(shape)
    quiet = shape.pop("quiet", False)
    with PhaseTransitionEngine(n, pods) as e:
        if quiet:
            cols = c_oracle.generate(SEED + 9, n, pods, **shape)
            cols["phase"][:] = po.PH_RUNNING                           # only ADDED-after-DELETED and tombstones get out
            e.ingest(cols)
            want, ost, dig = c_oracle.replay(cols, pods_hint=pods)
        else:
            e.generate(SEED + 9, n, pods, **shape)
            ost, dig = c_oracle.generate_replay(SEED + 9, n, pods, **shape)
        for env in ("staging", "production"):
            flt = FILTERS[env]
            e.set_filters(flt["ns_allow"], critical_only=flt["critical_only"])
            st = e.run()
            assert st["n_diff_rounds"] == {"deferred": 1, "looped": 2, "four_slices": 4}[mode]
            if env == "staging":
                assert {k: st[k] for k in ost} == ost and e.digest() == dig
                if quiet:
                    assert np.array_equal(e.fetch(columns=("idx",)).idx.astype(np.uint64), want)
            else:
                cols = e.read_log()
                want_p, ost_p, dig_p = c_oracle.replay(cols, critical_only=True, ns_allow=flt["ns_allow"], pods_hint=pods)
                assert {k: st[k] for k in ost_p} == ost_p and e.digest() == dig_p
                assert np.array_equal(e.fetch(columns=("idx",)).idx.astype(np.uint64), want_p)


def test_carried_batches_with_deferred_slices(monkeypatch):
    monkeypatch.setenv("PTE_DIFF_ROUND_MB", "1")
    n, pods = 24_000_000, 200_000
    cols = c_oracle.generate(SEED + 10, n, pods, dist=1, p_del=0.05, p_nostatus=0.02)
    want, ost, _ = c_oracle.replay(cols, pods_hint=pods)
    got, tomb = [], 0
    with PhaseTransitionEngine(n // 2, pods) as e:
        for lo, hi in ((0, n // 2), (n // 2, n)):
            e.ingest({k: v[lo:hi] for k, v in cols.items()})
            st = e.run(dedupe=True, carry=True)
            assert st["n_diff_rounds"] == 1 and st["n_launches"] >= 11
            got.append(e.fetch(columns=("idx",)).idx.astype(np.int64) + lo)
            tomb += st["n_tombstones"]
    assert np.array_equal(np.concatenate(got), want.astype(np.int64)) and tomb == ost["n_tombstones"]


# ---- errors are loud -------------------------------------------------------------
def test_bad_event_is_rejected(eng):
    cols = c_oracle.generate(7, 1000, 10)
    cols["type"][500] = 3          # an ERROR/BOOKMARK event: the reference would crash at pod_watcher.py:216
    eng.set_filters(None)
    eng.ingest(cols)
    with pytest.raises(PteError) as ei:
        eng.run()
    assert ei.value.code == -6


def test_zero_uid_is_rejected(eng):
    cols = c_oracle.generate(8, 1000, 10)
    cols["uid_hi"][3] = 0
    cols["uid_lo"][3] = 0
    eng.ingest(cols)
    with pytest.raises(PteError) as ei:
        eng.run()
    assert ei.value.code == -6


@pytest.mark.parametrize("env", list(FILTERS))
@pytest.mark.parametrize("dedupe", [True, False])
def test_unordered_log_replays_in_rv_order(eng, env, dedupe):
    """Replay order is ascending rv (ties: arrival order); a shuffled log gives the same notifications."""
    n, pods = 150_000, 2_000
    cols = c_oracle.generate(SEED + 9, n, pods, dist=1, p_del=0.07, p_nostatus=0.03, p_null=0.03)
    rng = np.random.default_rng(4)
    perm = rng.permutation(n)
    shuf = {k: v[perm] for k, v in cols.items()}
    st = check_against_oracle(eng, shuf, FILTERS[env], dedupe, pods)      # the oracle sorts by (rv, arrival) itself
    assert st["ordered"] == 0
    # ... and in terms of the ordered log: the same events, in the same order
    got = eng.fetch(columns=("idx",)).idx.astype(np.int64)
    want, _, _ = c_oracle.replay(cols, critical_only=FILTERS[env]["critical_only"], ns_allow=FILTERS[env]["ns_allow"],
                                 dedupe=dedupe, pods_hint=pods)
    assert np.array_equal(perm[got], want.astype(np.int64))
    # the engine's copy of the log is untouched
    back = eng.read_log()
    assert all(np.array_equal(back[k], shuf[k]) for k in shuf)


def test_unordered_log_with_rv_ties_and_offset(eng):
    n = 40_000
    cols = c_oracle.generate(21, n, 300, p_del=0.1)
    cols["rv"] = (cols["rv"] // np.uint64(7)) + np.uint64(1 << 40)       # ties, far from zero
    perm = np.random.default_rng(8).permutation(n)
    shuf = {k: v[perm] for k, v in cols.items()}
    check_against_oracle(eng, shuf, FILTERS["staging"], True, 300)


def test_unordered_log_with_too_wide_rv_span_is_rejected(eng):
    cols = c_oracle.generate(9, 5000, 10)
    cols["rv"][100] = np.uint64(1) << np.uint64(63)
    eng.set_filters(None)
    eng.ingest(cols)
    with pytest.raises(PteError) as ei:
        eng.run()
    assert ei.value.code == -7 and "span" in str(ei.value)


def test_unordered_stream_batch_is_rejected(eng):
    cols = c_oracle.generate(9, 5000, 10)
    cols["rv"][[100, 4000]] = cols["rv"][[4000, 100]]
    eng.ingest(cols)
    with pytest.raises(PteError) as ei:
        eng.run(carry=True)
    assert ei.value.code == -7


def test_too_many_pods_is_rejected():
    with PhaseTransitionEngine(100_000, 64) as e:     # table sized for 64 uids
        e.ingest(c_oracle.generate(10, 100_000, 50_000))
        with pytest.raises(PteError) as ei:
            e.run()
        assert ei.value.code == -5


def test_capacity_errors(eng):
    with pytest.raises(PteError):
        eng.ingest(c_oracle.generate(1, 4_000_001, 10))
    with pytest.raises(PteError):
        PhaseTransitionEngine(1 << 30, 10)
    fresh = PhaseTransitionEngine(1000, 10)
    with pytest.raises(PteError):
        fresh.fetch()
    fresh.close()


# ---- through the reference-shaped surface, against what the real reference produced ------
def _load(path):
    with gzip.open(path, "rb") as f:
        fx = json.loads(f.read())
    dt = dict(uid_hi=np.uint64, uid_lo=np.uint64, rv=np.uint64, ns_id=np.uint32, phase=np.uint8, type=np.uint8)
    fx["cols"] = {k: np.array(v, dtype=dt[k]) for k, v in fx["columns"].items()}
    return fx


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[4:-8] for p in GOLDEN])
def test_podwatcher_replay_matches_reference_payloads(path, monkeypatch):
    fx = _load(path)
    monkeypatch.chdir(ROOT)
    w = PodWatcher(fx["env"])
    assert w.config == fx["config"]
    log = ReplayLog.from_events(fk.events_from_columns(fx["cols"]))
    # R-trace: byte-identical wire bodies to the real reference's payloads
    batch = w.replay(log, dedupe=False)
    assert batch.indices.tolist() == fx["r_trace"]
    got = list(batch.payloads(event_timestamp=fx["frozen_timestamp"]))
    assert got == fx["payloads"]
    assert [json.dumps(d, allow_nan=False) for d in got] == [json.dumps(d, allow_nan=False) for d in fx["payloads"]]
    # T-trace: the transition layer applied to the real reference's payloads
    keep = po.transition_filter(fx["payloads"])
    tb = w.replay(log, dedupe=True)
    assert tb.indices.tolist() == [fx["r_trace"][k] for k in keep]
    assert list(tb.payloads(event_timestamp=fx["frozen_timestamp"])) == [fx["payloads"][k] for k in keep]


def test_live_loop_surface_with_mock_sink(monkeypatch):
    """start_watching() over a replayed stream (BASELINE config 1 shape) ends with the sink holding
    exactly the transition payloads, in order."""
    from k8s_watcher_b200 import ClusterApiClient
    from mock_transport import MockTransport
    fx = _load([p for p in GOLDEN if "c1_staging" in p][0])
    monkeypatch.chdir(ROOT)
    events = fk.events_from_columns(fx["cols"])
    fk.set_stream(events)
    sink = ClusterApiClient("http://clusterapi.invalid", transport=MockTransport())
    w = PodWatcher("staging", sink=sink)
    w.start_watching()
    keep = po.transition_filter(fx["payloads"])
    bodies = [json.loads(b) for b in sink.transport.bodies()]
    assert len(bodies) == len(keep)
    for b, k in zip(bodies, keep):
        want = dict(fx["payloads"][k])
        b.pop("event_timestamp"), want.pop("event_timestamp")       # wall clock, pod_watcher.py:199
        assert b == want


# ---- streams: state carried on the device from batch to batch ---------------------
@pytest.mark.parametrize("env", list(FILTERS))
@pytest.mark.parametrize("cuts", [(0.5,), (0.1, 0.11, 0.6), (0.0, 0.3, 0.3, 1.0)])
def test_carried_batches_equal_one_replay(env, cuts):
    n, pods = 120_000, 900
    cols = c_oracle.generate(SEED + 5, n, pods, dist=1, p_del=0.08, p_nostatus=0.03, p_null=0.03)
    flt = FILTERS[env]
    want, ost, _ = c_oracle.replay(cols, critical_only=flt["critical_only"], ns_allow=flt["ns_allow"], pods_hint=pods)
    edges = [0] + [int(c * n) for c in cuts] + [n]
    got, tomb = [], 0
    with PhaseTransitionEngine(n, pods) as e:
        e.set_filters(flt["ns_allow"], critical_only=flt["critical_only"])
        for lo, hi in zip(edges[:-1], edges[1:]):
            e.ingest({k: v[lo:hi] for k, v in cols.items()})
            st = e.run(dedupe=True, carry=True)
            got.append(e.fetch(columns=("idx",)).idx.astype(np.int64) + lo)
            tomb += st["n_tombstones"]
        assert np.array_equal(np.concatenate(got), want.astype(np.int64))
        assert tomb == ost["n_tombstones"] and st["n_pods"] <= pods
        # a run without the flag ends the stream: the same last batch replayed alone starts from scratch
        alone, _, _ = c_oracle.replay({k: v[lo:hi] for k, v in cols.items()}, critical_only=flt["critical_only"],
                                      ns_allow=flt["ns_allow"])
        e.run(dedupe=True)
        assert np.array_equal(e.fetch(columns=("idx",)).idx.astype(np.uint64), alone)
        # ... and the next carry run opens a NEW stream (no leftovers from the old one)
        e.run(dedupe=True, carry=True)
        assert np.array_equal(e.fetch(columns=("idx",)).idx.astype(np.uint64), alone)


def test_stream_batches_must_not_go_back_in_time():
    cols = c_oracle.generate(3, 2000, 50)
    with PhaseTransitionEngine(2000, 50) as e:
        e.ingest({k: v[1000:] for k, v in cols.items()})
        e.run(carry=True)
        e.ingest({k: v[:1000] for k, v in cols.items()})
        with pytest.raises(PteError) as ei:
            e.run(carry=True)
        assert ei.value.code == -7


def test_live_loop_micro_batches_match_one_replay(monkeypatch):
    """start_watching() with flush_every: the sink sees exactly the transitions of the whole stream."""
    from k8s_watcher_b200 import ClusterApiClient
    from mock_transport import MockTransport
    fx = _load([p for p in GOLDEN if "edge_development" in p][0])
    monkeypatch.chdir(ROOT)
    events = fk.events_from_columns(fx["cols"])
    fk.set_stream(events)
    sink = ClusterApiClient("http://clusterapi.invalid", transport=MockTransport())
    w = PodWatcher("staging", sink=sink, flush_every=64, flush_interval=None, max_pods=64, keep_batches=None)
    cfg = dict(fx["config"])
    cfg.pop("kubernetes", None)                                    # no kubeconfig file here: default set-up path
    w.environment, w.config = "development", cfg                   # development's namespace filter
    w.start_watching()
    assert len(w.batches) == -(-len(events) // 64)
    keep = po.transition_filter(fx["payloads"])
    bodies = [json.loads(b) for b in sink.transport.bodies()]
    assert [b["uid"] for b in bodies] == [fx["payloads"][k]["uid"] for k in keep]
    for b, k in zip(bodies, keep):
        want = dict(fx["payloads"][k])
        b.pop("event_timestamp"), want.pop("event_timestamp")
        assert b == want
    # a watcher that only keeps its last two batches alive notifies exactly the same
    fk.set_stream(events)
    sink2 = ClusterApiClient("http://clusterapi.invalid", transport=MockTransport())
    w2 = PodWatcher("staging", sink=sink2, flush_every=64, flush_interval=None, max_pods=64, keep_batches=2)
    w2.environment, w2.config = "development", cfg
    w2.start_watching()
    assert len(w2.batches) == 2
    assert [json.loads(x)["uid"] for x in sink2.transport.bodies()] == [x["uid"] for x in bodies]


# ---- payload serialisation on the device ------------------------------------------
def test_device_wire_bodies_match_requests_bytes(monkeypatch):
    from k8s_watcher_b200 import ClusterApiClient
    from k8s_watcher_b200.clusterapi_client import wire_body
    monkeypatch.chdir(ROOT)
    log = ReplayLog.synthetic(SEED, 60_000, 700, dist=1, p_del=0.1, p_nostatus=0.05, p_null=0.05)
    log.namespaces = ['de"fault', "kube\\system", "prodüction", "monitoring", "other"]
    sink = ClusterApiClient("http://clusterapi.invalid", transport=MockTransport())
    w = PodWatcher("production", sink=sink)          # production filters: namespace list + critical-only
    w.config["watcher"]["namespaces"] = ['de"fault', "prodüction", "monitoring", "kube\\system"]
    batch = w.replay(log)
    ts = "2025-06-01T12:00:00"
    got = batch.wire_bodies(event_timestamp=ts)
    want = [wire_body(p) for p in batch.payloads(event_timestamp=ts)]
    assert len(got) == len(want) == len(batch) > 1000
    assert got == want
    # the sink takes the whole batch in one call and sees the same bytes
    assert sink.update_pod_status_batch(batch, event_timestamp=ts) == len(batch)
    assert sink.transport.bodies() == want
    # a sub-range, and an empty batch
    from k8s_watcher_b200.pod_watcher import payload_fragments
    frags, n_phase, n_ns = payload_fragments(w, log, ts)
    buf, off = w._engine.serialize_payloads(frags, n_phase, n_ns, first=1500, count=1100)
    raw = buf.tobytes()
    assert [raw[int(off[i]):int(off[i + 1])] for i in range(1100)] == want[1500:2600]
    empty = w.replay(ReplayLog.synthetic(1, 0, 1))
    assert empty.wire_bodies(event_timestamp=ts) == []


# ---- raw watch-JSON ingestion on the device ----------------------------------------
def _json_case(n, pods, seed, **gen):
    cols = c_oracle.generate(seed, n, pods, **gen)
    events = fk.events_from_columns(cols)
    # a few pods in phases outside the well-known five, and one in a namespace that needs no escaping but is new
    for k in range(0, n, 97):
        if events[k]["object"].status is not None:
            events[k]["object"].status.phase = ("Terminating", "Evicted", "ContainerCreating")[k % 3]
    return events, fk.watch_json_lines(events, seed=seed)


@pytest.mark.parametrize("n,pods,seed", [(1, 1, 1), (2_000, 60, 2), (60_000, 900, 3)])
def test_watch_json_parses_to_the_same_columns_as_the_host_interner(n, pods, seed):
    events, text = _json_case(n, pods, seed, dist=1, p_del=0.08, p_nostatus=0.05, p_null=0.05)
    want = ReplayLog.from_events(events)
    with PhaseTransitionEngine(max(n, 1024), pods) as e:
        got = ReplayLog.from_watch_json(e, text)
    assert len(got) == n
    for k in ("uid_hi", "uid_lo", "rv", "type"):
        assert np.array_equal(got.columns[k], want.columns[k]), k
    # ids are handed out in order of first occurrence on both sides, so even the codes agree
    assert got.namespaces == want.namespaces and np.array_equal(got.columns["ns_id"], want.columns["ns_id"])
    assert got.phases == want.phases and np.array_equal(got.columns["phase"], want.columns["phase"])


@pytest.mark.parametrize("env", ["development", "staging", "production"])
def test_replay_of_watch_json_matches_reference_semantics(env, monkeypatch):
    monkeypatch.chdir(ROOT)
    events, text = _json_case(20_000, 300, 11, dist=0, p_del=0.1, p_nostatus=0.05, p_null=0.05)
    w = PodWatcher(env)
    batch = w.replay_watch_json(text)
    # oracle: the reference's filters on the very same pod objects, then the transition layer
    r = po.reference_trace(env, w.config, [(e["type"], e["object"]) for e in events], "T")
    keep = po.transition_filter(r)
    reach = [i for i, e in enumerate(events) if po.reaches_post_site(env, w.config, e["type"], e["object"])]
    assert batch.indices.tolist() == [reach[k] for k in keep]


@pytest.mark.parametrize("bad,reason", [
    (b'{"type":"ADDED","object":{"metadata":{"uid":5,"namespace":"a"}}}', "uid"),
    (b'{"type":"ADDED","object":{"metadata":{"namespace":"a"}}}', "uid"),
    (b'{"type":"BOOKMARK","object":{"metadata":{"uid":"00000000-0000-0000-0000-000000000001"}}}', "type"),
    (b'{"type":"ADDED","object":{"metadata":{"uid":"00000000-0000-0000-0000-000000000001","resourceVersion":"12x"}}}', "resourceVersion"),
    (b'{"type":"ADDED","object":{"metadata":{"uid":"00000000-0000-0000-0000-000000000001"}', "malformed"),
    (b'{"type":"ADDED","object":{"metadata":{"uid":"00000000-0000-0000-0000-000000000001","namespace":"a\\q"}}}', "escape"),
    (b'{"type":"ADDED","object":{"metadata":{"uid":"00000000-0000-0000-0000-000000000001","namespace":"\\ud83dx"}}}', "escape"),
])
def test_watch_json_bad_lines_are_named(bad, reason):
    good = b'{"type":"ADDED","object":{"metadata":{"uid":"00000000-0000-0000-0000-00000000000a","namespace":"x","resourceVersion":"5"},"status":{"phase":"Running"}}}'
    with PhaseTransitionEngine(1024, 16) as e:
        with pytest.raises(PteError) as ei:
            e.ingest_json(good + b"\n" + good + b"\n" + bad + b"\n" + good + b"\n")
        assert ei.value.code == -10 and "line 2" in str(ei.value) and reason in str(ei.value)
        assert e.ingest_json(good + b"\n" + good) == 2          # and the engine is still usable
        assert e.json_strings(0) == ["x"]
