"""Randomised parity: the CUDA path against the C oracle on logs hypothesis makes up.  Needs a B200: -m gpu.

The shapes the fixed tests pin down one by one (ragged sizes, hot pods, tombstones, status-less pods,
filters, rv disorder with ties, stream batches) are mixed freely here; every example is checked record
by record.  Examples are small so that a few hundred of them fit in seconds.
"""
import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings, strategies as st

from k8s_watcher_b200 import PhaseTransitionEngine
from oracle import c_oracle

pytestmark = pytest.mark.gpu

MAX_N = 20000
COMMON = dict(deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow],
              derandomize=True)


@pytest.fixture(scope="module")
def eng():
    e = PhaseTransitionEngine(MAX_N, MAX_N)
    yield e
    e.close()


log_shapes = st.fixed_dictionaries(dict(
    seed=st.integers(0, 2**63), n=st.integers(0, MAX_N), pods=st.integers(1, 400), dist=st.integers(0, 1),
    p_del=st.sampled_from([0.0, 0.02, 0.3, 1.0]), p_nostatus=st.sampled_from([0.0, 0.05, 0.5]),
    p_null=st.sampled_from([0.0, 0.05, 0.5])))
filters = st.fixed_dictionaries(dict(
    critical_only=st.booleans(),
    ns_allow=st.one_of(st.none(), st.lists(st.integers(0, 6), min_size=0, max_size=5, unique=True)),
    ns_filter_on=st.booleans()))


def make_log(shape):
    s = dict(shape)
    return c_oracle.generate(s.pop("seed"), s.pop("n"), s.pop("pods"), **s)


def oracle_allow(flt):
    """The oracle takes one allow-list; an active filter with no names lets nothing through."""
    if flt["ns_allow"]:
        return flt["ns_allow"]
    return [10**6] if (flt["ns_filter_on"] and flt["ns_allow"] is not None) else None


def set_filters(eng, flt):
    eng.set_filters(flt["ns_allow"], ns_filter_on=flt["ns_filter_on"] and flt["ns_allow"] is not None,
                    critical_only=flt["critical_only"])


@settings(max_examples=400, **COMMON)
@given(shape=log_shapes, flt=filters, dedupe=st.booleans())
def test_random_logs_match_the_oracle(eng, shape, flt, dedupe):
    cols = make_log(shape)
    set_filters(eng, flt)
    eng.ingest(cols)
    stt = eng.run(dedupe=dedupe)
    got = eng.fetch()
    want, ost, dig = c_oracle.replay(cols, critical_only=flt["critical_only"], ns_allow=oracle_allow(flt), dedupe=dedupe,
                                     pods_hint=shape["pods"])
    assert {k: stt[k] for k in ost} == ost
    assert np.array_equal(got.idx.astype(np.uint64), want)
    i = want.astype(np.int64)
    assert np.array_equal(got.rv, cols["rv"][i]) and np.array_equal(got.ns_id, cols["ns_id"][i])
    assert np.array_equal(got.phase, cols["phase"][i]) and np.array_equal(got.type, cols["type"][i])
    assert eng.digest() == dig


@settings(max_examples=150, **COMMON)
@given(shape=log_shapes, flt=filters, perm_seed=st.integers(0, 2**32 - 1), ties=st.booleans())
def test_random_unordered_logs_replay_in_rv_order(eng, shape, flt, perm_seed, ties):
    cols = make_log(shape)
    n = shape["n"]
    rng = np.random.default_rng(perm_seed)
    if ties and n:
        cols["rv"] = (cols["rv"] // np.uint64(3)) + np.uint64(7)            # runs of equal resourceVersions
    order = rng.permutation(n)                                              # arrival order != rv order
    shuffled = {k: v[order] for k, v in cols.items()}
    set_filters(eng, flt)
    eng.ingest(shuffled)
    stt = eng.run(dedupe=True)
    got = eng.fetch(columns=("idx", "rv"))
    # the oracle replays the log sorted by (rv, arrival); its indices point into that sorted log
    by_rv = np.argsort(shuffled["rv"], kind="stable")
    want, ost, _ = c_oracle.replay({k: v[by_rv] for k, v in shuffled.items()}, critical_only=flt["critical_only"],
                                   ns_allow=oracle_allow(flt), pods_hint=shape["pods"])
    assert {k: stt[k] for k in ost} == ost
    assert np.array_equal(got.idx.astype(np.int64), by_rv[want.astype(np.int64)])
    assert stt["ordered"] == (1 if n < 2 or bool(np.all(np.diff(shuffled["rv"].astype(np.int64)) >= 0)) else 0)


@settings(max_examples=150, **COMMON)
@given(shape=log_shapes, flt=filters, cuts=st.lists(st.floats(0.0, 1.0), min_size=1, max_size=5))
def test_random_stream_batches_add_up_to_one_replay(shape, flt, cuts):
    cols = make_log(shape)
    n = shape["n"]
    want, ost, _ = c_oracle.replay(cols, critical_only=flt["critical_only"], ns_allow=oracle_allow(flt), pods_hint=shape["pods"])
    edges = [0] + sorted(int(c * n) for c in cuts) + [n]
    got, tomb = [], 0
    with PhaseTransitionEngine(max(n, 1), MAX_N) as e:
        set_filters(e, flt)
        for lo, hi in zip(edges[:-1], edges[1:]):
            e.ingest({k: v[lo:hi] for k, v in cols.items()})
            stt = e.run(dedupe=True, carry=True)
            got.append(e.fetch(columns=("idx",)).idx.astype(np.int64) + lo)
            tomb += stt["n_tombstones"]
    assert np.array_equal(np.concatenate(got), want.astype(np.int64))
    assert tomb == ost["n_tombstones"]


# ---- the raw watch text as a stream, cut at random --------------------------------------------------------------
json_shapes = st.fixed_dictionaries(dict(
    seed=st.integers(0, 2**31), n=st.integers(1, 500), pods=st.integers(1, 40), dist=st.integers(0, 1),
    p_del=st.sampled_from([0.0, 0.1, 0.6]), p_nostatus=st.sampled_from([0.0, 0.1]), p_null=st.sampled_from([0.0, 0.1])))


@settings(max_examples=150, **COMMON)
@given(shape=json_shapes, env=st.sampled_from(["development", "staging", "production"]), cut_seed=st.integers(0, 2**31),
       dedupe=st.booleans())
def test_random_watch_text_in_random_chunks(shape, env, cut_seed, dedupe):
    """PodWatcher.watch_json on made-up streams: whatever the chunking, trailing newline or not, the notifications are the
    reference's filters followed by the transition layer (dedupe) or the reference's own trace (no dedupe)."""
    import os
    import fake_kubernetes as fk
    from k8s_watcher_b200 import PodWatcher
    from oracle import podwatch_oracle as po
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cols = make_log(shape)
    n = len(cols["rv"])
    events = fk.events_from_columns(cols)
    lines = [l for l in fk.watch_json_lines(events, seed=shape["seed"] % 1000).split(b"\n") if l]
    where = {l: i for i, l in enumerate(lines)}
    rnd = np.random.default_rng(cut_seed)
    cuts = [0]
    while cuts[-1] < n:
        cuts.append(min(n, cuts[-1] + int(rnd.choice([1, 2, 3, 17, 64, 200]))))
    chunks = [b"\n".join(lines[a:b]) + (b"\n" if rnd.random() < 0.5 else b"") for a, b in zip(cuts[:-1], cuts[1:])]
    chunks.insert(int(rnd.integers(0, len(chunks) + 1)), b"")               # an empty read changes nothing
    offs = iter(cuts[:-1])
    w = PodWatcher(env, config_dir=os.path.join(root, "config"), max_pods=256, keep_batches=None)
    w.logger.disabled = True
    batches = w.watch_json(iter(chunks), lambda line: events[where[line]]["object"], dedupe=dedupe)
    got = [int(i) + a for b, a in zip(batches, offs) for i in b.indices.tolist()]
    reach = [i for i, e in enumerate(events) if po.reaches_post_site(env, w.config, e["type"], e["object"])]
    if dedupe:
        r = po.reference_trace(env, w.config, [(e["type"], e["object"]) for e in events], "T")
        want = [reach[k] for k in po.transition_filter(r)]
    else:
        want = reach
    assert got == want
    w._engine.close()
