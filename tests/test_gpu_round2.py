"""Round-2 parity tests (need a B200: -m gpu): full-size digests, per-shard replays, the emit bitmap, streams
whose pods churn far beyond max_pods, the partition-first path, and direct numpy / oracle anchors for the
device JSON parser and the device-built request bodies."""
import json
import os

import numpy as np
import pytest

import fake_kubernetes as fk
from k8s_watcher_b200 import PhaseTransitionEngine, PodWatcher, PteError, ReplayLog, generate_host
from oracle import c_oracle
from oracle import podwatch_oracle as po

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SEED = 0x6B38735F77617463
with open(os.path.join(ROOT, "tests", "golden", "digests.json")) as _f:
    DIGESTS = json.load(_f)
FILTERS = {
    "staging": dict(critical_only=False, ns_allow=None),
    "development": dict(critical_only=False, ns_allow=[0, 1]),
    "production": dict(critical_only=True, ns_allow=[0, 1, 2, 3]),
}


def _row(name, G, g):
    return DIGESTS["rows"]["%s/%d/%d" % (name, G, g)]


def _check_row(eng, st, row):
    assert (st["n_in"], st["n_survived"], st["n_emitted"], st["n_tombstones"]) == (
        row["n_in"], row["n_survived"], row["n_emitted"], row["n_tombstones"])
    assert "%016x" % eng.digest() == row["digest"]


# ---- full size: BASELINE configs 3 and 4 at 1 B events, by count + order-sensitive digest -------------------
@pytest.mark.parametrize("name", ["c3", "c4"])
def test_one_billion_events_match_the_oracle_digest(name):
    w = DIGESTS["workloads"][name]
    with PhaseTransitionEngine(w["events_per_shard"], w["pods_per_shard"]) as e:
        e.generate(SEED, w["events_per_shard"], w["pods_per_shard"], dist=w["dist"], p_del=w["p_del"])
        _check_row(e, e.run(), _row(name, 1, 0))
        # idempotence at full size: a second replay of the resident log gives the same records
        _check_row(e, e.run(), _row(name, 1, 0))


# ---- uid-hash shards on the GPU: every shard equals the oracle's replay of that shard, and the shards add up ----
def test_two_gpu_shards_equal_the_oracle_shard_by_shard():
    n_per, pods_per, G = 400_000, 6_000, 2
    kw = dict(dist=1, p_del=0.05, p_nostatus=0.02)
    parts, tot = [], np.zeros(4, np.int64)
    with PhaseTransitionEngine(n_per, pods_per) as e:
        for g in range(G):
            e.generate(SEED, n_per, pods_per, shard=g, n_shards=G, **kw)
            st = e.run()
            got = e.fetch(columns=("idx",))
            cols = c_oracle.generate(SEED, n_per, pods_per, shard=g, n_shards=G, **kw)
            idx, ost, dig = c_oracle.replay(cols, pods_hint=pods_per)
            assert np.array_equal(got.idx.astype(np.uint64), idx) and e.digest() == dig
            assert {k: st[k] for k in ost} == ost
            tot += [st["n_in"], st["n_survived"], st["n_emitted"], st["n_tombstones"]]
            parts.append(cols)
    allc = {k: np.concatenate([p[k] for p in parts]) for k in parts[0]}
    order = np.argsort(allc["rv"], kind="stable")
    _, st_all, _ = c_oracle.replay({k: v[order] for k, v in allc.items()}, pods_hint=pods_per * G)
    assert [int(x) for x in tot] == [st_all[k] for k in ("n_in", "n_survived", "n_emitted", "n_tombstones")]


def test_config2_shards_match_the_digest_table():
    w = DIGESTS["workloads"]["c2"]
    with PhaseTransitionEngine(w["events_per_shard"], w["pods_per_shard"]) as e:
        for G in (1, 2):
            for g in range(G):
                e.generate(SEED, w["events_per_shard"], w["pods_per_shard"], dist=w["dist"], p_del=w["p_del"], shard=g, n_shards=G)
                _check_row(e, e.run(), _row("c2", G, g))


# ---- the emit bitmap says what fetch(idx) says; the namespace column may stay at home -------------------------
@pytest.mark.parametrize("dedupe", [True, False])
def test_emit_bitmap_and_log_without_namespace_column(dedupe):
    n, pods = 300_001, 4_000
    cols = c_oracle.generate(SEED + 1, n, pods, dist=1, p_del=0.05, p_null=0.02)
    want, ost, _ = c_oracle.replay(cols, dedupe=dedupe, pods_hint=pods)
    with PhaseTransitionEngine(n, pods) as e:
        e.ingest(cols, with_namespaces=False)
        st = e.run(dedupe=dedupe)
        words = e.fetch_bitmap()
        bits = np.unpackbits(words.view(np.uint8), bitorder="little")[:n]
        assert np.array_equal(np.flatnonzero(bits).astype(np.uint64), want) and st["n_emitted"] == ost["n_emitted"]
        got = e.fetch(columns=("idx", "ns_id"))
        assert np.array_equal(got.idx.astype(np.uint64), want) and not got.ns_id.any()
        # a namespace list needs the column
        e.set_filters([0, 1])
        with pytest.raises(PteError) as ei:
            e.ingest(cols, with_namespaces=False)
        assert ei.value.code == -1
        e.ingest(cols)
        e.run(dedupe=dedupe)
        want2, _, _ = c_oracle.replay(cols, ns_allow=[0, 1], dedupe=dedupe, pods_hint=pods)
        bits = np.unpackbits(e.fetch_bitmap().view(np.uint8), bitorder="little")[:n]
        assert np.array_equal(np.flatnonzero(bits).astype(np.uint64), want2)


def test_emit_bitmap_is_refused_for_a_presorted_log():
    cols = c_oracle.generate(SEED, 5000, 100)
    perm = np.random.default_rng(1).permutation(5000)
    with PhaseTransitionEngine(5000, 100) as e:
        e.ingest({k: v[perm] for k, v in cols.items()})
        assert e.run()["ordered"] == 0
        with pytest.raises(PteError) as ei:
            e.fetch_bitmap()
        assert ei.value.code == -8


# ---- a stream whose pods churn far beyond max_pods: the dictionary is rebuilt from its live pods ---------------
def _churn_columns(max_pods, generations, seed=3):
    """`generations` waves of max_pods/2 pods; every pod lives ADDED .. DELETED inside its wave, so at most
    max_pods/2 are alive at any time while generations * max_pods / 2 distinct uids pass through."""
    rng = np.random.default_rng(seed)
    per = max_pods // 2
    hi, lo, ph, ty = [], [], [], []
    for g in range(generations):
        uh = rng.integers(1, 1 << 63, per, dtype=np.uint64)
        ul = rng.integers(1, 1 << 63, per, dtype=np.uint64)
        life = rng.integers(2, 7, per)                       # events per pod, the last one DELETED
        owner = np.repeat(np.arange(per), life)
        rng.shuffle(owner)                                   # interleave the pods of the wave, keep per-pod order
        seen = np.zeros(per, np.int64)
        for p in owner:
            k = seen[p]
            seen[p] += 1
            hi.append(uh[p]); lo.append(ul[p])
            ty.append(0 if k == 0 else (2 if k == life[p] - 1 else 1))
            ph.append(int(rng.integers(0, 3)))
    n = len(hi)
    return {"uid_hi": np.array(hi, np.uint64), "uid_lo": np.array(lo, np.uint64), "rv": np.arange(1, n + 1, dtype=np.uint64),
            "ns_id": np.zeros(n, np.uint32), "phase": np.array(ph, np.uint8), "type": np.array(ty, np.uint8)}


def test_stream_with_fifty_times_max_pods_distinct_uids(monkeypatch):
    monkeypatch.chdir(ROOT)
    max_pods = 1000
    cols = _churn_columns(max_pods, generations=100)          # 50 x max_pods distinct uids
    n = len(cols["rv"])
    assert len(set(zip(cols["uid_hi"].tolist(), cols["uid_lo"].tolist()))) == 50 * max_pods
    want, ost, _ = c_oracle.replay(cols, pods_hint=1 << 16)
    # engine level, batches of 777 events
    got, rebuilds = [], 0
    with PhaseTransitionEngine(1024, max_pods) as e:
        for lo in range(0, n, 777):
            e.ingest({k: v[lo:lo + 777] for k, v in cols.items()})
            st = e.run(carry=True)
            got.append(e.fetch(columns=("idx",)).idx.astype(np.int64) + lo)
            rebuilds = st["n_rebuilds"]
        assert np.array_equal(np.concatenate(got), want.astype(np.int64))
        assert rebuilds >= 10 and st["ms_rebuild"] > 0
        print("stream of %d events, %d distinct uids, max_pods %d: %d dictionary rebuilds, last one %.3f ms" % (
            n, 50 * max_pods, max_pods, rebuilds, st["ms_rebuild"]))
    # the same through the watcher's live surface: flush_every far below the number of distinct uids
    w = PodWatcher("staging", flush_every=512, flush_interval=None, max_pods=max_pods, keep_batches=None)
    for ev in fk.events_from_columns({k: v[:60_000] for k, v in cols.items()}, rich=False):
        w.handle_pod_event(ev["type"], ev["object"])
    w.flush()
    idx, base = [], 0
    for b in w.batches:
        idx.append(b.indices.astype(np.int64) + base)
        base += len(b.log)
    want60, _, _ = c_oracle.replay({k: v[:60_000] for k, v in cols.items()}, pods_hint=1 << 16)
    assert np.array_equal(np.concatenate(idx), want60.astype(np.int64))


def test_stream_orders_by_arrival_not_by_resource_version(monkeypatch):
    """A list+watch hands over its initial ADDED burst in key order, with resourceVersions in any order, and the
    reference never looks at them: the live loop must take the events as they come."""
    monkeypatch.chdir(ROOT)
    cols = c_oracle.generate(SEED + 9, 3000, 80, dist=1, p_del=0.1)
    events = fk.events_from_columns(cols, rich=False)
    rng = np.random.default_rng(2)
    for ev in events:                                          # scrambled, partly missing resource versions
        ev["object"].metadata.resource_version = None if rng.random() < 0.3 else str(int(rng.integers(1, 1000)))
    for dedupe in (True, False):
        w = PodWatcher("staging", dedupe=dedupe, flush_every=256, flush_interval=None, keep_batches=None)
        for ev in events:
            w.handle_pod_event(ev["type"], ev["object"])
        w.flush()
        idx, base = [], 0
        for b in w.batches:
            idx.append(b.indices.astype(np.int64) + base)
            base += len(b.log)
        want, _, _ = c_oracle.replay(cols, dedupe=dedupe, pods_hint=80)
        assert np.array_equal(np.concatenate(idx), want.astype(np.int64))


def test_engine_grows_in_place_and_keeps_the_stream():
    cols = c_oracle.generate(SEED + 2, 40_000, 300, dist=1, p_del=0.05)
    want, _, _ = c_oracle.replay(cols, pods_hint=300)
    with PhaseTransitionEngine(1000, 300) as e:
        got, lo = [], 0
        for size in (1000, 9000, 30_000):                     # each batch longer than the engine was built for
            if size > e.max_events:
                e.grow(size)
            e.ingest({k: v[lo:lo + size] for k, v in cols.items()})
            e.run(carry=True)
            got.append(e.fetch(columns=("idx",)).idx.astype(np.int64) + lo)
            lo += size
        assert np.array_equal(np.concatenate(got), want.astype(np.int64))


# ---- the partition-first path (PTE_PATH=partition) is held to the same oracle ----------------------------------
@pytest.mark.parametrize("env", list(FILTERS))
@pytest.mark.parametrize("shape", [
    dict(n=1000, pods=50, dist=0, p_del=0.05),                               # BASELINE config 1 shape
    dict(n=700_001, pods=9_000, dist=0, p_del=0.0, p_nostatus=0.02, p_null=0.02),
    dict(n=3_000_000, pods=40_000, dist=1, p_del=0.05),                      # skew: cut partitions, cross-segment fix-up
    dict(n=2_000_000, pods=3, dist=1, p_del=0.3),                            # three pods: every tile is one hot slot
])
def test_partition_path_matches_oracle(monkeypatch, env, shape):
    monkeypatch.setenv("PTE_PATH", "partition")
    kw = dict(shape)
    n, pods = kw.pop("n"), kw.pop("pods")
    flt = FILTERS[env]
    cols = c_oracle.generate(SEED + 7, n, pods, **kw)
    idx, ost, dig = c_oracle.replay(cols, critical_only=flt["critical_only"], ns_allow=flt["ns_allow"], pods_hint=pods)
    with PhaseTransitionEngine(n, max(pods, 5000)) as e:
        e.set_filters(flt["ns_allow"], critical_only=flt["critical_only"])
        e.ingest(cols)
        st = e.run()
        assert st["path"] == 1
        got = e.fetch()
        assert {k: st[k] for k in ost} == ost
        assert np.array_equal(got.idx.astype(np.uint64), idx) and e.digest() == dig
        i = idx.astype(np.int64)
        assert np.array_equal(got.type, cols["type"][i]) and np.array_equal(got.rv, cols["rv"][i])
        if len(i):                                               # the key is a bijection of the uid on this path too
            uid = cols["uid_hi"][i].astype(object) * (1 << 64) + cols["uid_lo"][i].astype(object)
            assert len(set(zip(uid.tolist(), got.key.tolist()))) == len(set(uid.tolist())) == len(set(got.key.tolist()))
        # and it agrees with the sort path record by record
        monkeypatch.setenv("PTE_PATH", "sort")
    with PhaseTransitionEngine(n, max(pods, 5000)) as e2:
        e2.set_filters(flt["ns_allow"], critical_only=flt["critical_only"])
        e2.ingest(cols)
        assert e2.run()["path"] == 0
        assert np.array_equal(e2.fetch(columns=("idx",)).idx, got.idx)


def test_partition_path_unordered_log_and_adversarial_uids(monkeypatch):
    monkeypatch.setenv("PTE_PATH", "partition")
    n, pods = 200_000, 2_000
    cols = c_oracle.generate(SEED + 8, n, pods, dist=1, p_del=0.05)
    # uids that share one half, or have a zero half: partition and dictionary must still tell them apart
    cols["uid_lo"][::3] = cols["uid_lo"][0]
    cols["uid_hi"][1::3] = 0
    perm = np.random.default_rng(4).permutation(n)
    shuffled = {k: v[perm] for k, v in cols.items()}
    want, ost, dig = c_oracle.replay(shuffled, pods_hint=n)
    with PhaseTransitionEngine(n, n) as e:
        e.ingest(shuffled)
        st = e.run()
        assert st["ordered"] == 0 and st["n_emitted"] == ost["n_emitted"]
        assert np.array_equal(e.fetch(columns=("idx",)).idx.astype(np.uint64), want)


# ---- direct anchors: device JSON columns and device-built bodies against oracle / numpy ------------
def test_device_json_columns_against_the_source_columns():
    """The columns the device parser extracts from watch-JSON text equal the integer columns the text was made from
    (no product code on the expected side)."""
    n, pods = 30_000, 700
    cols = c_oracle.generate(SEED + 3, n, pods, dist=1, p_del=0.05, p_nostatus=0.03, p_null=0.03)
    events = fk.events_from_columns(cols)
    text = fk.watch_json_lines(events, seed=5)
    with PhaseTransitionEngine(n, pods) as e:
        assert e.ingest_json(text) == n
        dev = e.read_log()
        names = e.json_strings(0)
    for k in ("uid_hi", "uid_lo", "type"):
        assert np.array_equal(dev[k], cols[k]), k
    assert np.array_equal(dev["rv"], np.array([int(ev["object"].metadata.resource_version) for ev in events], np.uint64))
    assert [names[i] for i in dev["ns_id"]] == [fk.NAMESPACES[i] for i in cols["ns_id"]]
    # phases: the five well-known ones and the two "absent" codes are fixed by include/pte.h
    assert np.array_equal(dev["phase"], cols["phase"])


def test_device_wire_bodies_against_json_dumps_of_oracle_payloads(monkeypatch):
    monkeypatch.chdir(ROOT)
    n, pods = 20_000, 300
    cols = c_oracle.generate(SEED + 4, n, pods, dist=1, p_del=0.05, p_nostatus=0.03, p_null=0.03)
    log = ReplayLog(cols, list(fk.NAMESPACES))
    w = PodWatcher("staging")
    batch = w.replay(log)
    ts = "2025-01-01T00:00:00"
    bodies = batch.wire_bodies(event_timestamp=ts)
    # expected side: the oracle's restatement of _extract_pod_data on the same minimal pods, serialised like requests does
    idx, _, _ = c_oracle.replay(cols, pods_hint=pods)
    assert len(bodies) == len(idx)
    for b, i in zip(bodies[:4000], idx[:4000].astype(np.int64)):
        pod = log.pod_at(int(i))
        want = po.extract_pod_data("staging", pod, ts)
        want["event_type"] = po.TYPE_NAMES[int(cols["type"][i])]
        assert b == json.dumps(want, allow_nan=False).encode("utf-8")


# ---- pipelined replay: copies overlap the replay chunk by chunk, the result is the emit bitmap -------------------
@pytest.mark.parametrize("env", list(FILTERS))
@pytest.mark.parametrize("dedupe", [True, False])
def test_pipelined_replay_equals_one_replay(env, dedupe):
    n, pods = 250_003, 3_000
    flt = FILTERS[env]
    cols = c_oracle.generate(SEED + 6, n, pods, dist=1, p_del=0.06, p_nostatus=0.02, p_null=0.02)
    want, ost, _ = c_oracle.replay(cols, critical_only=flt["critical_only"], ns_allow=flt["ns_allow"], dedupe=dedupe, pods_hint=pods)
    with PhaseTransitionEngine(n, pods) as e:
        e.set_filters(flt["ns_allow"], critical_only=flt["critical_only"])
        for chunk in (8192, 65_536, 1 << 20):                    # 31 chunks, 4 chunks, one chunk
            words, st = e.replay_pipelined(cols, dedupe=dedupe, chunk_events=chunk, with_namespaces=flt["ns_allow"] is not None)
            bits = np.unpackbits(words.view(np.uint8), bitorder="little")[:n]
            assert np.array_equal(np.flatnonzero(bits).astype(np.uint64), want), chunk
            assert {k: st[k] for k in ost} == ost
        with pytest.raises(PteError) as ei:                      # the columns were not assembled
            e.fetch()
        assert ei.value.code == -8
        # a log that is not in rv order is refused here (the plain path pre-sorts it)
        perm = np.random.default_rng(0).permutation(n)
        with pytest.raises(PteError) as ei:
            e.replay_pipelined({k: v[perm] for k, v in cols.items()}, dedupe=dedupe, chunk_events=65_536)
        assert ei.value.code == -7


def test_podwatcher_replay_pipelines_long_logs(monkeypatch):
    monkeypatch.chdir(ROOT)
    n, pods = 300_000, 2_000
    cols = c_oracle.generate(SEED + 10, n, pods, dist=0, p_del=0.05)
    log = ReplayLog.pinned(n, list(fk.NAMESPACES))
    for k, v in cols.items():
        log.columns[k][:] = v
    try:
        for envname in ("staging", "development"):
            flt = FILTERS[envname]
            w = PodWatcher(envname)
            w.pipeline_min_events, w.pipeline_chunk_events = 1000, 65_536
            batch = w.replay(log)
            assert batch.bitmap is not None and batch.stats["n_passes"] > 3          # several chunks went through the sort
            want, ost, _ = c_oracle.replay(cols, ns_allow=flt["ns_allow"], pods_hint=pods)
            assert len(batch) == ost["n_emitted"]
            assert np.array_equal(batch.indices.astype(np.uint64), want)
            assert np.array_equal(batch.emitted.type, cols["type"][want.astype(np.int64)])
            # the shuffled log falls back to the path that pre-sorts
            perm = np.random.default_rng(3).permutation(n)
            sh = ReplayLog({k: v[perm] for k, v in cols.items()}, list(fk.NAMESPACES))
            b2 = w.replay(sh)
            want2, _, _ = c_oracle.replay(sh.columns, ns_allow=flt["ns_allow"], pods_hint=pods)
            assert b2.bitmap is None and np.array_equal(b2.indices.astype(np.uint64), want2)
    finally:
        log.free()


def test_pipelined_replay_without_rv_column_for_arrival_ordered_logs():
    """A log declared to be in arrival order needs no resourceVersion column on the device: positions are the order."""
    n, pods = 150_001, 2_000
    cols = c_oracle.generate(SEED + 12, n, pods, dist=1, p_del=0.06)
    want, ost, _ = c_oracle.replay(cols, pods_hint=pods)
    junk = dict(cols, rv=np.zeros(n, np.uint64))                   # must not be looked at
    with PhaseTransitionEngine(n, pods) as e:
        words, st = e.replay_pipelined(junk, chunk_events=32_768, with_namespaces=False, with_rv=False)
        bits = np.unpackbits(words.view(np.uint8), bitorder="little")[:n]
        assert np.array_equal(np.flatnonzero(bits).astype(np.uint64), want) and st["n_emitted"] == ost["n_emitted"]


# ---- watch JSON spelled with escape sequences: strings are what they stand for ------------------
def _respell(text: str, rnd) -> str:
    """The JSON string body of `text` with some characters written as \\uXXXX (or the short escapes)."""
    out = []
    for ch in text:
        r = rnd.random()
        cp = ord(ch)
        if cp > 0xFFFF:
            hi, lo = 0xD800 + ((cp - 0x10000) >> 10), 0xDC00 + ((cp - 0x10000) & 0x3FF)
            out.append("\\u%04x\\u%04X" % (hi, lo) if r < 0.5 else ch)
        elif ch in '"\\' or cp < 0x20:
            out.append(json.dumps(ch)[1:-1])
        elif r < 0.25:
            out.append("\\u%04x" % cp if rnd.random() < 0.5 else "\\u%04X" % cp)
        elif ch == "/" and r < 0.6:
            out.append("\\/")
        else:
            out.append(ch)
    return "".join(out)


@pytest.mark.gpu
def test_device_json_decodes_escapes_in_every_field_it_reads():
    import random
    rnd = random.Random(99)
    n, pods = 6000, 150
    cols = c_oracle.generate(SEED, n, pods, dist=1, p_del=0.1, p_nostatus=0.05, p_null=0.05)
    events = fk.events_from_columns(cols)
    odd_ns = ['quo"te', "back\\slash", "sl/ash", "é中\U0001f600", "tab\there", "plain-but-new"]
    odd_ph = ["Terminating", 'Weird"Phase', "Évicted\U0001f4a5"]
    for k, ev in enumerate(events):
        if k % 5 == 0:
            ev["object"].metadata.namespace = odd_ns[(k // 5) % len(odd_ns)]
        if k % 7 == 0 and ev["object"].status is not None:
            ev["object"].status.phase = odd_ph[(k // 7) % len(odd_ph)]
    lines = fk.watch_json_lines(events, seed=4).split(b"\n")
    lines = [l for l in lines if l]
    assert len(lines) == n
    out = []
    for ev, line in zip(events, lines):
        text = line.decode()
        md, st = ev["object"].metadata, ev["object"].status
        vals = [ev["type"], md.uid, md.namespace, md.resource_version] + ([st.phase] if st is not None and st.phase is not None else [])
        for v in vals:
            for spelled in {json.dumps(v), json.dumps(v, ensure_ascii=False)}:
                text = text.replace(spelled, '"' + _respell(v, rnd) + '"')
        assert json.loads(text) == json.loads(line)          # same document, other spelling
        out.append(text.encode())
    text = b"\n".join(out) + b"\n"
    want = ReplayLog.from_events(events)
    with PhaseTransitionEngine(n, pods) as e:
        got = ReplayLog.from_watch_json(e, text)
    for k in ("uid_hi", "uid_lo", "rv", "type"):
        assert np.array_equal(got.columns[k], want.columns[k]), k
    assert got.namespaces == want.namespaces and np.array_equal(got.columns["ns_id"], want.columns["ns_id"])
    assert got.phases == want.phases and np.array_equal(got.columns["phase"], want.columns["phase"])
    # ... and against what the text was made from, without any product code in between
    for k in ("uid_hi", "uid_lo", "rv", "type"):
        assert np.array_equal(got.columns[k], cols[k]), k
    assert [got.namespaces[i] for i in got.columns["ns_id"].tolist()] == [ev["object"].metadata.namespace for ev in events]
    for code, ev in zip(got.columns["phase"].tolist(), events):
        st = ev["object"].status
        assert (code == 6) if st is None else (code == 5) if st.phase is None else got.phases[code] == st.phase


# ---- the raw watch stream as a live loop (SURVEY 8f rows 1 + 2): chunks of lines, state carried on the device ---------
@pytest.mark.gpu
@pytest.mark.parametrize("env", ["development", "staging", "production"])
def test_raw_watch_stream_in_chunks_notifies_like_the_object_loop(env, monkeypatch):
    """PodWatcher.watch_json: the same notifications as the reference's filters + transition layer over the whole
    stream, however the stream is cut into chunks; pod objects are made only for lines that lead to a notification."""
    from k8s_watcher_b200 import ClusterApiClient
    from mock_transport import MockTransport
    monkeypatch.chdir(ROOT)
    n, pods = 30_000, 250
    cols = c_oracle.generate(SEED + 21, n, pods, dist=1, p_del=0.1, p_nostatus=0.05, p_null=0.05)
    events = fk.events_from_columns(cols)
    odd = ("Terminating", "Evicted", "ContainerCreating", "Weird\\Phase")
    for k in range(0, n, 61):
        if events[k]["object"].status is not None:
            events[k]["object"].status.phase = odd[(k // 61 * 7 + k) % len(odd)]     # custom phases, first seen in different chunks
    for k in range(0, n, 97):
        events[k]["object"].metadata.namespace = ("late-namespace-%d" % (k % 5), "production")[k % 2]
    lines = [l for l in fk.watch_json_lines(events, seed=8).split(b"\n") if l]
    assert len(lines) == n
    where = {l: i for i, l in enumerate(lines)}
    assert len(where) == n
    rnd = np.random.default_rng(3)
    cuts = [0]
    while cuts[-1] < n:
        cuts.append(min(n, cuts[-1] + int(rnd.choice([1, 5, 300, 2500, 4096, 4097, 7000]))))
    chunks = [b"\n".join(lines[a:b]) + (b"\n" if k % 2 else b"") for k, (a, b) in enumerate(zip(cuts[:-1], cuts[1:]))]
    calls = []

    def deserialize(line):
        calls.append(where[line])
        return events[where[line]]["object"]

    sink = ClusterApiClient("http://clusterapi.invalid", transport=MockTransport())
    w = PodWatcher(env, sink=sink, max_pods=1024, keep_batches=None)
    batches = w.watch_json(iter(chunks), deserialize)
    got = np.concatenate([b.indices.astype(np.int64) + a for b, a in zip(batches, cuts[:-1])])
    r = po.reference_trace(env, w.config, [(e["type"], e["object"]) for e in events], "T")
    keep = po.transition_filter(r)
    reach = [i for i, e in enumerate(events) if po.reaches_post_site(env, w.config, e["type"], e["object"])]
    want = [reach[k] for k in keep]
    assert got.tolist() == want
    assert sorted(calls) == want                                   # objects only for what is notified
    bodies = [json.loads(b) for b in sink.transport.bodies()]
    assert [b["uid"] for b in bodies] == [events[i]["object"].metadata.uid for i in want]
    assert [b["event_type"] for b in bodies] == [events[i]["type"] for i in want]
    # the stream mode ends with the loop: a one-shot replay of the whole text afterwards is its own log again
    one = w.replay_watch_json(b"\n".join(lines))
    assert one.indices.tolist() == want


@pytest.mark.gpu
def test_device_json_keeps_uid_strings_of_any_form_apart(monkeypatch):
    """metadata.uid is whatever string arrives (pod_watcher.py:164): canonical UUIDs are their own key, every other string
    gets a key of the reserved block -- one string one key, also across the chunks of a stream, and the replay equals the
    object loop's."""
    monkeypatch.chdir(ROOT)
    n, pods = 12_000, 300
    cols = c_oracle.generate(SEED + 31, n, pods, dist=1, p_del=0.1, p_nostatus=0.03, p_null=0.03)
    events = fk.events_from_columns(cols)
    canon = {}
    for ev in events:                      # every third pod gets a uid that is no canonical UUID, in several flavours
        md = ev["object"].metadata
        p = int(md.name.split("-")[1])
        if p % 3 == 0:
            flavours = ("pod-uid-%d", "%032x", "ABCDEF00-0000-4000-8000-%012X", "ffffffff-ffff-ffff-0000-%012x", "wei\\rd \"uid\" é %d")
            canon.setdefault(md.uid, flavours[p // 3 % len(flavours)] % p)
            md.uid = canon[md.uid]
    events[7]["object"].metadata.uid = ""                                     # even the empty string is a uid string
    events[9]["object"].metadata.uid = "00000000-0000-0000-0000-000000000000"   # the all-zero uid cannot be its own key
    text = fk.watch_json_lines(events, seed=12)
    lines = [l for l in text.split(b"\n") if l]
    with PhaseTransitionEngine(n, pods + 10) as e:
        log = ReplayLog.from_watch_json(e, text)
    keys = list(zip(log.columns["uid_hi"].tolist(), log.columns["uid_lo"].tolist()))
    strings = [ev["object"].metadata.uid for ev in events]
    by_key, by_string = {}, {}
    for k, s_ in zip(keys, strings):
        assert by_key.setdefault(k, s_) == s_ and by_string.setdefault(s_, k) == k          # a bijection
    reserved = [(k, s_) for k, s_ in by_key.items() if k[0] == 0xFFFFFFFFFFFFFFFF]
    assert len(reserved) == len({s_ for s_ in strings if by_string[s_][0] == 0xFFFFFFFFFFFFFFFF}) > 90
    for k, s_ in by_key.items():
        assert log.uids.string(*k) == s_ if k[0] == 0xFFFFFFFFFFFFFFFF else codes_uid(k) == s_
    # and the whole thing as a stream in chunks: same notifications as the object loop with the same odd uids
    where = {l: i for i, l in enumerate(lines)}
    w = PodWatcher("staging", max_pods=1024, keep_batches=None)
    cuts = list(range(0, n, 1700)) + [n]
    chunks = [b"\n".join(lines[a:b]) for a, b in zip(cuts[:-1], cuts[1:])]
    batches = w.watch_json(iter(chunks), lambda line: events[where[line]]["object"])
    got = np.concatenate([b.indices.astype(np.int64) + a for b, a in zip(batches, cuts[:-1])])
    r = po.reference_trace("staging", w.config, [(ev["type"], ev["object"]) for ev in events], "T")
    assert got.tolist() == po.transition_filter(r)


def codes_uid(k):
    from k8s_watcher_b200 import codes
    return codes.uid_to_str(*k)


@pytest.mark.gpu
def test_raw_watch_loop_after_an_object_loop_on_the_same_watcher(monkeypatch):
    """watch_json starts its own stream even if the watcher's engine carried another one before, and leaves the engine
    ready for one-shot logs when it ends -- also when it ends with an error."""
    monkeypatch.chdir(ROOT)
    n, pods = 5000, 60
    cols = c_oracle.generate(SEED + 41, n, pods, dist=0, p_del=0.1)
    events = fk.events_from_columns(cols)
    lines = [l for l in fk.watch_json_lines(events, seed=1).split(b"\n") if l]
    where = {l: i for i, l in enumerate(lines)}
    r = po.transition_filter(po.reference_trace("staging", PodWatcher("staging").config, [(e["type"], e["object"]) for e in events], "T"))
    w = PodWatcher("staging", flush_every=1000, flush_interval=None, max_pods=256, keep_batches=None)
    for ev in events[:2500]:                                   # an object loop first: its stream is at arrival number 2500
        w.handle_pod_event(ev["type"], ev["object"])
    w.flush()
    chunks = [b"\n".join(lines[a:a + 900]) for a in range(0, n, 900)]
    batches = w.watch_json(iter(chunks), lambda line: events[where[line]]["object"])
    got = np.concatenate([b.indices.astype(np.int64) + 900 * k for k, b in enumerate(batches)])
    assert got.tolist() == r
    with pytest.raises(PteError):                              # a broken line ends the loop ...
        w.watch_json(iter([chunks[0], b'{"type":"ADDED"', chunks[1]]), lambda line: events[where[line]]["object"])
    one = w.replay_watch_json(b"\n".join(lines))               # ... and the engine takes one-shot logs as before
    assert one.indices.tolist() == r


@pytest.mark.gpu
def test_pipelined_replay_with_the_packed_type_and_phase_byte(monkeypatch):
    """type << 6 | phase in one byte per event instead of two columns (17 B/event over the wire): same notifications."""
    from k8s_watcher_b200.engine import pack_meta
    monkeypatch.chdir(ROOT)
    n, pods = 200_000, 1500
    cols = c_oracle.generate(SEED + 51, n, pods, dist=1, p_del=0.07, p_nostatus=0.04, p_null=0.04)
    cols["phase"][::13] = 62                                                   # the largest phase code still fits in six bits
    want, ost, _ = c_oracle.replay(cols, pods_hint=pods)
    meta = pack_meta(cols["type"], cols["phase"])
    assert np.array_equal(meta >> 6, cols["type"]) and np.array_equal(meta & 63, cols["phase"])
    with PhaseTransitionEngine(n, pods) as e:
        for chunk in (8192, 65_536, 0):
            words, st = e.replay_pipelined(cols, chunk_events=chunk, with_rv=False, with_namespaces=False, meta=meta)
            bits = np.unpackbits(words.view(np.uint8), bitorder="little")[:n]
            assert np.array_equal(np.nonzero(bits)[0].astype(np.uint64), want) and st["n_tombstones"] == ost["n_tombstones"]
            back = e.read_log()
            assert np.array_equal(back["phase"], cols["phase"]) and np.array_equal(back["type"], cols["type"])
    # through the watcher: a pinned log that keeps the packed byte
    log = ReplayLog.pinned(n, list(fk.NAMESPACES))
    for k, v in cols.items():
        log.columns[k][:] = v
    log.arrival_ordered = True
    log.pack_meta()
    try:
        w = PodWatcher("staging")
        w.pipeline_min_events = 1000
        batch = w.replay(log)
        assert batch.bitmap is not None and batch.indices.tolist() == want.tolist()
    finally:
        log.free()
