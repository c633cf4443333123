"""Host-side logic and the C-ABI surface; no GPU needed (and no device compute is attempted)."""
import json
import os
import re

import numpy as np
import pytest

import fake_kubernetes as fk
from k8s_watcher_b200 import (ClusterApiClient, PhaseTransitionEngine, PodWatcher, PteError,
                              ReplayLog, _abi, codes, generate_host, shard_of)
from k8s_watcher_b200.clusterapi_client import ConnectionFailed, RequestTimeout, wire_body
from mock_transport import MockTransport
from oracle import c_oracle
from oracle import podwatch_oracle as po

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "pte.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(pte_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(_abi.SYMBOLS), declared ^ set(_abi.SYMBOLS)
    L = _abi.lib()
    for name in declared:
        assert getattr(L, name) is not None


def test_struct_layouts_match_header(tmp_path):
    """sizeof/offsetof as gcc sees include/pte.h == what the ctypes mirror uses."""
    import ctypes as C
    import subprocess
    src = tmp_path / "layout.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "pte.h"\n'
                   'int main(void){printf("%zu %zu %zu %zu %zu\\n", sizeof(pte_stats), offsetof(pte_stats, ms_total),'
                   ' offsetof(pte_stats, n_passes), sizeof(pte_gen_params), offsetof(pte_gen_params, shard));return 0;}\n')
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(src)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()]
    assert got == [C.sizeof(_abi.PteStats), _abi.PteStats.ms_total.offset, _abi.PteStats.n_passes.offset,
                   C.sizeof(_abi.PteGenParams), _abi.PteGenParams.shard.offset]


def test_codes_agree_with_oracle():
    assert codes.TYPE_NAMES == po.TYPE_NAMES and codes.PHASE_NAMES == po.PHASE_NAMES
    assert (codes.PH_NULL, codes.PH_NOSTATUS) == (po.PH_NULL, po.PH_NOSTATUS)


@pytest.mark.parametrize("kw", [dict(dist=0, p_del=0.05), dict(dist=1, p_del=0.1, p_nostatus=0.05, p_null=0.05),
                                dict(dist=0, shard=2, n_shards=4), dict(dist=1, shard=7, n_shards=8, p_del=0.05)])
def test_host_generator_matches_oracle_generator(kw):
    a = generate_host(42, 50_000, 700, **kw)
    b = c_oracle.generate(42, 50_000, 700, **kw)
    for k in b:
        assert np.array_equal(a[k], b[k]), k


def test_shard_of_matches_oracle_and_generator():
    L = c_oracle.lib()
    cols = generate_host(1, 2000, 300, shard=5, n_shards=8)
    for hi, lo in zip(cols["uid_hi"][:200], cols["uid_lo"][:200]):
        assert shard_of(int(hi), int(lo), 8) == 5 == L.orc_shard_of(int(hi), 8)


def test_no_device_means_loud_failure():
    if _abi.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(PteError) as ei:
        PhaseTransitionEngine(1000, 10)
    assert ei.value.code == -3
    w = PodWatcher("staging", config_dir=os.path.join(ROOT, "config"))
    with pytest.raises(PteError):
        w.replay(ReplayLog.synthetic(1, 100, 5))


def test_uid_string_round_trip():
    for hi, lo in ((1, 2), (0xFFFFFFFFFFFFFFFF, 0), (0x0123456789ABCDEF, 0xFEDCBA9876543210)):
        assert codes.str_to_uid(codes.uid_to_str(hi, lo)) == (hi, lo)
    assert codes.uid_to_str(0x0123456789ABCDEF, 0xFEDCBA9876543210) == "01234567-89ab-cdef-fedc-ba9876543210"
    with pytest.raises(ValueError):
        codes.str_to_uid("not-a-uuid")


def test_uid_interner_is_exact_for_any_string():
    """metadata.uid is whatever string arrives (pod_watcher.py:164): different strings never share a key, one string
    always has the same key, canonical UUIDs keep their own value."""
    u = codes.UidInterner()
    canon = "01234567-89ab-cdef-fedc-ba9876543210"
    assert u.key(canon) == (0x0123456789ABCDEF, 0xFEDCBA9876543210) and len(u) == 0
    odd = ["not-a-uuid", "", canon.upper(), canon.replace("-", ""), "0123456-789ab-cdef-fedc-ba9876543210", canon + " ",
           "ffffffff-ffff-ffff-0000-000000000001", "00000000-0000-0000-0000-000000000000", "pod/with/slashes", "é"]
    keys = [u.key(s) for s in odd]
    assert len(set(keys)) == len(odd) and all(hi == codes.UidInterner.RESERVED_HI and lo > 0 for hi, lo in keys)
    assert [u.key(s) for s in odd] == keys and len(u) == len(odd)
    assert [u.string(*k) for k in keys] == odd and u.string(*u.key(canon)) == canon
    # the reserved block is not reachable by a canonical string: the one that would land there goes through the dictionary
    assert u.key("ffffffff-ffff-ffff-0000-000000000002")[1] == len(odd) + 1


def test_from_events_keeps_non_uuid_uids_apart_across_batches():
    uids = codes.UidInterner()
    names = []
    a = [{"type": "ADDED", "object": fk.make_pod(1, "weird uid A", "default", "Running")},
         {"type": "ADDED", "object": fk.make_pod(2, "weird uid B", "default", "Running")}]
    b = [{"type": "MODIFIED", "object": fk.make_pod(2, "weird uid B", "default", "Failed")},
         {"type": "ADDED", "object": fk.make_pod(3, codes.uid_to_str(7, 9), "default", "Running")}]
    la = ReplayLog.from_events(a, namespaces=names, uids=uids)
    lb = ReplayLog.from_events(b, namespaces=names, uids=uids)
    ka = list(zip(la.columns["uid_hi"].tolist(), la.columns["uid_lo"].tolist()))
    kb = list(zip(lb.columns["uid_hi"].tolist(), lb.columns["uid_lo"].tolist()))
    assert ka[0] != ka[1] and kb[0] == ka[1] and kb[1] == (7, 9)
    assert lb.uids.string(*kb[0]) == "weird uid B"


def test_replay_log_from_events_round_trips_columns():
    cols = c_oracle.generate(3, 500, 20, dist=1, p_del=0.1, p_nostatus=0.1, p_null=0.1)
    log = ReplayLog.from_events(fk.events_from_columns(cols))
    for k in ("uid_hi", "uid_lo", "rv", "phase", "type"):
        assert np.array_equal(log.columns[k], cols[k]), k
    # namespaces are interned in first-seen order; names must map back
    names = [log.namespaces[i] for i in log.columns["ns_id"]]
    assert names == [fk.NAMESPACES[i] for i in cols["ns_id"]]
    assert log.namespace_ids(["default", "nope"]) == [log.namespaces.index("default")]


def test_replay_log_interns_unknown_phase_strings_and_rejects_bad_types():
    pod = fk.make_pod(1, codes.uid_to_str(5, 6), "default", "Terminating")
    log = ReplayLog.from_events([{"type": "ADDED", "object": pod}])
    assert log.columns["phase"][0] == codes.PH_FIRST_DYNAMIC and log.phases[7] == "Terminating"
    assert log.pod_at(0) is pod
    with pytest.raises(ValueError):
        ReplayLog.from_events([{"type": "ERROR", "object": pod}])


def test_synthetic_log_pods_and_sharding():
    log = ReplayLog.synthetic(9, 3000, 100, p_del=0.05, p_nostatus=0.1)
    pod = log.pod_at(0)
    assert pod.metadata.namespace in codes.DEFAULT_NAMESPACES and len(pod.metadata.uid) == 36
    i = int(np.nonzero(log.columns["phase"] == codes.PH_NOSTATUS)[0][0])
    assert log.pod_at(i).status is None
    parts = log.shard(4)
    assert sum(len(p) for p in parts) == len(log)
    for g, p in enumerate(parts):
        assert (np.diff(p.columns["rv"].astype(np.int64)) > 0).all()
        assert all(shard_of(int(h), int(l), 4) == g for h, l in zip(p.columns["uid_hi"][:50], p.columns["uid_lo"][:50]))
    # sharded replay == unsharded replay restricted to the shard (oracle, the property multi-GPU rests on)
    idx_all, _, _ = c_oracle.replay(log.columns)
    owner = np.array([shard_of(int(h), int(l), 4) for h, l in zip(log.columns["uid_hi"], log.columns["uid_lo"])])
    for g, p in enumerate(parts):
        local, _, _ = c_oracle.replay(p.columns)
        pos = np.nonzero(owner == g)[0]
        assert np.array_equal(pos[local.astype(np.int64)], idx_all[owner[idx_all.astype(np.int64)] == g].astype(np.int64))


# ---- PodWatcher surface (reference: watcher/pod_watcher.py) ---------------------------------
@pytest.fixture
def in_root(monkeypatch):
    monkeypatch.chdir(ROOT)


def test_config_regimes_match_reference(in_root):
    dev, stg, prod = PodWatcher("development"), PodWatcher("staging"), PodWatcher("production")
    assert dev.config["watcher"]["namespaces"] == ["default", "kube-system"]
    assert dev.config["watcher"]["log_level"] == "DEBUG" and dev.config["watcher"]["retry"]["max_attempts"] == 3
    assert "namespaces" not in stg.config["watcher"] and stg.config["clusterapi"]["timeout"] == 30
    assert prod.config["watcher"]["alerts"]["critical_events_only"] is True
    assert prod.config["clusterapi"]["auth"]["api_key"] == ""        # ${PROD_CLUSTERAPI_API_KEY} unset -> ""
    assert not dev._critical_only() and not stg._critical_only() and prod._critical_only()
    for w in (dev, stg, prod):
        assert w.v1 is None and hasattr(w.watch, "stream") and w.logger is not None


def test_env_substitution_and_missing_config(in_root, monkeypatch, tmp_path, capsys):
    monkeypatch.setenv("CLUSTERAPI_API_KEY", "s3cret")
    assert PodWatcher("development").config["clusterapi"]["auth"]["api_key"] == "s3cret"
    w = PodWatcher("staging")
    assert w._substitute_env_vars({"a": ["${NOPE_X:-dflt}", "x${NOPE}"]}) == {"a": ["dflt", "x${NOPE}"]}
    assert w._merge_configs({"a": {"b": 1, "c": [1]}, "d": 1}, {"a": {"c": [2]}, "e": 2}) == \
        {"a": {"b": 1, "c": [2]}, "d": 1, "e": 2}
    monkeypatch.chdir(tmp_path)
    assert PodWatcher("staging").config == {}           # missing files -> {} (pod_watcher.py:39-41)
    assert "not found" in capsys.readouterr().out


def test_per_event_surface_matches_oracle(in_root):
    cols = c_oracle.generate(5, 300, 10, p_del=0.1, p_nostatus=0.1, p_null=0.1)
    events = fk.events_from_columns(cols)
    for env in ("development", "staging", "production"):
        w = PodWatcher(env)
        for ev in events:
            assert w.should_process_event(ev["type"], ev["object"]) == \
                po.should_process_event(env, w.config, ev["type"], ev["object"])
            assert w._extract_pod_data(ev["object"], event_timestamp="T") == po.extract_pod_data(env, ev["object"], "T")
        d = w._extract_pod_data(events[0]["object"])
        assert list(d) == ["name", "namespace", "uid", "environment", "status", "spec", "metadata", "event_timestamp"]


def test_setup_failure_returns_false_and_start_watching_returns(in_root, caplog):
    w = PodWatcher("development")            # kubeconfig ./assets/config does not exist here
    assert w.setup_k8s_client() is False
    assert w.start_watching() is None        # logs and returns, never raises (pod_watcher.py:245-247)
    assert PodWatcher("staging").setup_k8s_client() is True


def test_stream_failure_is_logged_and_reraised(in_root):
    w = PodWatcher("staging")

    class Boom:
        stopped = False

        def stream(self, *a, **k):
            raise RuntimeError("stream broke")
            yield

        def stop(self):
            self.stopped = True
    w.watch = Boom()
    with pytest.raises(RuntimeError):
        w.start_watching()
    assert w.watch.stopped                   # finally: watch.stop() (pod_watcher.py:276-277)


# ---- sink (reference: watcher/clusterapi_client.py) --------------------------------------------
def test_sink_contract():
    t = MockTransport()
    c = ClusterApiClient("http://h:1/", "key", transport=t)
    assert c.base_url == "http://h:1" and c.headers["Authorization"] == "Bearer key"
    assert c.update_pod_status({"name": "p", "x": None}) is True
    assert t.requests[0]["url"] == "http://h:1/api/pods/update" and t.bodies() == [b'{"name": "p", "x": null}']
    assert ClusterApiClient("http://h", transport=MockTransport(500)).update_pod_status({}) is False
    assert c.health_check() is True and ClusterApiClient("http://h", transport=MockTransport(503)).health_check() is False

    def raising(exc):
        def send(*a):
            raise exc
        return send
    for exc in (ConnectionFailed("x"), RequestTimeout("x"), ValueError("x")):
        cl = ClusterApiClient("http://h", transport=raising(exc))
        assert cl.update_pod_status({"name": "p"}) is False and cl.health_check() is False
    assert c.update_pod_status({"bad": float("nan")}) is False       # allow_nan=False, never raises
    assert wire_body({"a": 1, "b": [1, {"c": None}]}) == json.dumps({"a": 1, "b": [1, {"c": None}]}).encode()


def test_payload_fragments_rebuild_the_reference_wire_body(in_root):
    """The fragments handed to pte_serialize_payloads, glued together the way the kernel does it, are
    byte-identical to json.dumps of the payload -- including names that need escaping."""
    from k8s_watcher_b200.pod_watcher import payload_fragments
    log = ReplayLog.synthetic(4, 400, 9, p_del=0.2, p_nostatus=0.2, p_null=0.2)
    log.namespaces = ['de"fault', "kube\\system", "prodüction", "monitoring\n", "other"]
    w = PodWatcher("staging")
    w.environment = 'st"agingé'
    frags, n_phase, n_ns = payload_fragments(w, log, "2025-01-01T00:00:00")
    segs, types, phases, nss = frags[:6], frags[6:9], frags[9:9 + n_phase], frags[9 + n_phase:]
    assert len(nss) == n_ns == 5 and n_phase >= 7
    for i in range(len(log)):
        c = log.columns
        uid = codes.uid_to_str(c["uid_hi"][i], c["uid_lo"][i]).encode()
        body = (segs[0] + uid[:13] + segs[1] + nss[int(c["ns_id"][i])] + segs[2] + uid + segs[3] +
                phases[int(c["phase"][i])] + segs[4] + types[int(c["type"][i])] + segs[5])
        want = w._extract_pod_data(log.pod_at(i), event_timestamp="2025-01-01T00:00:00")
        want["event_type"] = log.type_at(i)
        assert body == wire_body(want)


def test_sink_retry_timeout_and_config_factory(in_root):
    """The knobs config/base.yaml declares (clusterapi.timeout, clusterapi.retry.*) and the reference never reads."""
    calls, naps = [], []

    def flaky(method, url, body, headers, timeout):
        calls.append((method, timeout))
        return 200 if len(calls) >= 3 else 503
    w = PodWatcher("staging")
    c = ClusterApiClient.from_config(w.config, transport=flaky, sleep=naps.append)
    assert (c.base_url, c.timeout, c.max_attempts, c.delay_seconds) == ("http://localhost:3000", 30, 3, 2.0)
    assert c.update_pod_status({"name": "p"}) is True            # third attempt succeeds
    assert calls == [("POST", 30)] * 3 and naps == [2.0, 2.0]
    calls.clear(); naps.clear()
    always_down = ClusterApiClient("http://h", transport=MockTransport(500), max_attempts=2, delay_seconds=0)
    assert always_down.update_pod_status({"name": "p"}) is False and len(always_down.transport.requests) == 2
    # defaults keep the reference's behaviour: one attempt, no timeout
    one = ClusterApiClient("http://h", transport=MockTransport(500))
    assert one.update_pod_status({}) is False and len(one.transport.requests) == 1 and one.timeout is None
    prod = ClusterApiClient.from_config(PodWatcher("production").config)
    assert prod.base_url == "https://prod-clusterapi.example.com" and prod.api_key is None


def test_health_gate_warns_and_continues(in_root, caplog):
    import logging
    fk.set_stream([])
    w = PodWatcher("staging", sink=ClusterApiClient("http://h", transport=MockTransport(503)))
    with caplog.at_level(logging.WARNING):
        w.start_watching()                                         # empty stream: nothing to replay, no GPU needed
    assert any("health check failed" in r.message for r in caplog.records)
