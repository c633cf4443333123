"""The oracle against what the real reference produced (tests/golden, made by make_golden.py)."""
import glob
import gzip
import json
import os

import numpy as np
import pytest

import fake_kubernetes as fk
from oracle import c_oracle
from oracle import podwatch_oracle as po

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "ref_*.json.gz")))


def load(path):
    with gzip.open(path, "rb") as f:
        fx = json.loads(f.read())
    dt = dict(uid_hi=np.uint64, uid_lo=np.uint64, rv=np.uint64, ns_id=np.uint32, phase=np.uint8, type=np.uint8)
    fx["cols"] = {k: np.array(v, dtype=dt[k]) for k, v in fx["columns"].items()}
    return fx


def filter_inputs(fx):
    """(critical_only, ns_allow ids) the way the reference derives them from its config."""
    cfg, env = fx["config"], fx["env"]
    crit = env == "production" and bool(cfg.get("watcher", {}).get("alerts", {}).get("critical_events_only", False))
    names = cfg.get("watcher", {}).get("namespaces", []) or []
    allow = [fk.NAMESPACES.index(n) for n in names if n in fk.NAMESPACES]
    return crit, allow, bool(names)


def test_six_fixtures_present():
    assert len(GOLDEN) == 6


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[4:-8] for p in GOLDEN])
def test_generator_reproduces_fixture_columns(path):
    fx = load(path)
    g = fx["gen"]
    cols = c_oracle.generate(g["seed"], g["n"], g["pods"], dist=g["dist"], p_del=g["p_del"],
                             p_nostatus=g["p_nostatus"], p_null=g["p_null"])
    for k in cols:
        assert np.array_equal(cols[k], fx["cols"][k]), k


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[4:-8] for p in GOLDEN])
def test_python_restatement_matches_reference_payloads(path):
    fx = load(path)
    events = [(e["type"], e["object"]) for e in fk.events_from_columns(fx["cols"])]
    got = po.reference_trace(fx["env"], fx["config"], events, fx["frozen_timestamp"])
    assert got == fx["payloads"]
    # key order matters on the wire (json.dumps keeps insertion order)
    assert [list(d) for d in got] == [list(d) for d in fx["payloads"]]
    assert json.dumps(got) == json.dumps(fx["payloads"])


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[4:-8] for p in GOLDEN])
def test_column_replay_matches_reference_r_trace(path):
    fx = load(path)
    crit, allow, has_list = filter_inputs(fx)
    c = fx["cols"]
    idx_py, st_py = po.replay_columns(c["uid_hi"], c["uid_lo"], c["rv"], c["ns_id"], c["phase"], c["type"],
                                      critical_only=crit, ns_allow=allow, dedupe=False)
    assert idx_py == fx["r_trace"]
    idx_c, st_c, _ = c_oracle.replay(c, critical_only=crit, ns_allow=allow, dedupe=False)
    assert idx_c.tolist() == fx["r_trace"]
    assert st_c == st_py


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[4:-8] for p in GOLDEN])
def test_transition_layer_three_ways(path):
    """T-trace: spec on real payloads == column Python oracle == C oracle (layer itself is unpinned upstream)."""
    fx = load(path)
    crit, allow, _ = filter_inputs(fx)
    c = fx["cols"]
    keep = po.transition_filter(fx["payloads"])
    want = [fx["r_trace"][k] for k in keep]
    idx_py, st_py = po.replay_columns(c["uid_hi"], c["uid_lo"], c["rv"], c["ns_id"], c["phase"], c["type"],
                                      critical_only=crit, ns_allow=allow, dedupe=True)
    assert idx_py == want
    idx_c, st_c, dig_c = c_oracle.replay(c, critical_only=crit, ns_allow=allow, dedupe=True)
    assert idx_c.tolist() == want
    assert st_c == st_py
    assert dig_c == po.digest(want, c["uid_hi"], c["uid_lo"], c["rv"], c["ns_id"], c["phase"], c["type"])
    assert 0 < len(want) <= len(fx["r_trace"])


def test_edge_stream_has_the_edge_cases():
    fx = load([p for p in GOLDEN if "edge_staging" in p][0])
    ph = fx["cols"]["phase"]
    assert (ph == po.PH_NOSTATUS).any() and (ph == po.PH_NULL).any()
    phases = [p["status"]["phase"] for p in fx["payloads"]]
    assert None in phases and "Unknown" in phases
    types = fx["cols"]["type"]
    assert (types == po.TYPE_DELETED).any() and (types == po.TYPE_ADDED).sum() > 12  # re-ADD after delete


def test_production_keeps_statusless_and_drops_null_phase():
    """pod_watcher.py:209 edge cases the survey probed: status=None kept, phase=None dropped."""
    fx = load([p for p in GOLDEN if "edge_production" in p][0])
    c = fx["cols"]
    kept = set(fx["r_trace"])
    ns_ok = np.isin(c["ns_id"], [0, 1, 2, 3])
    for i in range(len(c["rv"])):
        if not ns_ok[i] or c["type"][i] == po.TYPE_DELETED:
            continue
        if c["phase"][i] == po.PH_NOSTATUS:
            assert i in kept
        if c["phase"][i] == po.PH_NULL:
            assert i not in kept


@pytest.mark.parametrize("dist,p_del", [(0, 0.0), (0, 0.05), (1, 0.05)])
def test_streaming_oracle_equals_materialised(dist, p_del):
    seed, n, pods = 12345, 200_000, 3_000
    cols = c_oracle.generate(seed, n, pods, dist=dist, p_del=p_del, p_nostatus=0.01, p_null=0.01)
    for crit, allow in ((False, None), (True, [0, 2, 3])):
        idx, st, dig = c_oracle.replay(cols, critical_only=crit, ns_allow=allow, dedupe=True, pods_hint=pods)
        st2, dig2 = c_oracle.generate_replay(seed, n, pods, dist=dist, p_del=p_del, p_nostatus=0.01, p_null=0.01,
                                             critical_only=crit, ns_allow=allow, dedupe=True)
        assert st == st2 and dig == dig2


def test_python_and_c_oracle_agree_on_random_logs():
    rng = np.random.default_rng(7)
    for trial in range(20):
        n = int(rng.integers(1, 400))
        pods = int(rng.integers(1, 30))
        cols = c_oracle.generate(int(rng.integers(1 << 62)), n, pods, dist=int(trial % 2), p_del=0.1,
                                 p_nostatus=0.05, p_null=0.05)
        crit = bool(trial % 3 == 0)
        allow = [0, 1] if trial % 4 == 0 else None
        for dedupe in (True, False):
            a, sa = po.replay_columns(cols["uid_hi"], cols["uid_lo"], cols["rv"], cols["ns_id"], cols["phase"],
                                      cols["type"], critical_only=crit, ns_allow=allow, dedupe=dedupe)
            b, sb, _ = c_oracle.replay(cols, critical_only=crit, ns_allow=allow, dedupe=dedupe)
            assert a == b.tolist() and sa == sb


def test_unordered_log_replays_in_rv_order():
    cols = c_oracle.generate(99, 500, 20, p_del=0.1)
    ordered, _, dig = c_oracle.replay(cols)
    perm = np.random.default_rng(3).permutation(500)
    shuf = {k: v[perm] for k, v in cols.items()}
    got, _, dig2 = c_oracle.replay(shuf)
    assert np.array_equal(perm[got.astype(np.int64)], ordered.astype(np.int64))
    assert dig2 == dig
    a, _ = po.replay_columns(shuf["uid_hi"], shuf["uid_lo"], shuf["rv"], shuf["ns_id"], shuf["phase"], shuf["type"])
    assert a == got.tolist()


def test_shard_function_inverts_uid_construction():
    L = c_oracle.lib()
    for G in (2, 4, 8):
        for g in range(G):
            cols = c_oracle.generate(5, 300, 40, shard=g, n_shards=G)
            assert all(L.orc_shard_of(int(h), G) == g for h in cols["uid_hi"])
            assert ((cols["rv"] - 1) % G == g).all()


def test_empty_log():
    cols = c_oracle.generate(1, 0, 10)
    idx, st, dig = c_oracle.replay(cols)
    assert idx.size == 0 and st["n_emitted"] == 0 and dig == 0
