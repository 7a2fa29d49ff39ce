"""TF_CONFIG resolver: roles, chief election, rank order, malformed input (SURVEY 4.3 tier 1)."""
import json

import pytest
from hypothesis import given, settings, strategies as st

from tfdl.cluster import TFConfigClusterResolver as R, TFConfigError, make_tf_config

REF = {"cluster": {"worker": ["172.16.16.5:12345", "172.16.16.6:12345"]}, "task": {"type": "worker", "index": 1}}


def test_reference_script_config():
    r = R.from_env(json.dumps(REF))  # tf_dist_example.py:6-10
    assert r.task_type == "worker" and r.task_id == 1 and r.num_workers == 2
    assert not r.is_chief and r.id_in_cluster == 1
    assert r.address == "172.16.16.6:12345" and r.leader_address == "172.16.16.5:12345"
    assert json.loads(r.to_tf_config()) == REF


def test_worker0_is_chief_without_chief_job():
    r = R.from_env({"cluster": REF["cluster"], "task": {"type": "worker", "index": 0}})
    assert r.is_chief and r.id_in_cluster == 0


def test_explicit_chief_takes_rank_zero():
    cfg = {"cluster": {"chief": ["a:1"], "worker": ["b:1", "c:1"], "ps": ["p:1"], "evaluator": ["e:1"]},
           "task": {"type": "worker", "index": 0}}
    r = R.from_env(cfg)
    assert not r.is_chief and r.id_in_cluster == 1 and r.num_workers == 3
    assert r.worker_addresses == ["a:1", "b:1", "c:1"]
    assert R.from_env({**cfg, "task": {"type": "chief", "index": 0}}).is_chief
    ps = R.from_env({**cfg, "task": {"type": "ps", "index": 0}})
    assert not ps.is_training_task
    with pytest.raises(TFConfigError):
        _ = ps.id_in_cluster


def test_env_and_override_and_missing(monkeypatch):
    monkeypatch.delenv("TF_CONFIG", raising=False)
    r = R.from_env()
    assert r.is_single_worker and r.is_chief and r.num_workers == 1 and r.id_in_cluster == 0
    monkeypatch.setenv("TF_CONFIG", make_tf_config(["h:1", "h:2"], "worker", 1))
    assert R.from_env().task_id == 1
    assert R.from_env(make_tf_config(["h:1", "h:2"], "worker", 0)).task_id == 0  # explicit override wins


@pytest.mark.parametrize("bad", [
    "{not json", '["list"]',
    {"cluster": {"worker": ["h:1"]}, "task": {"type": "worker", "index": 3}},
    {"cluster": {"worker": ["h:1"]}, "task": {"type": "boss", "index": 0}},
    {"cluster": {"worker": ["nocolon"]}, "task": {"type": "worker", "index": 0}},
    {"cluster": {"worker": ["h:99999"]}, "task": {"type": "worker", "index": 0}},
    {"cluster": {"worker": ["h:1", "h:1"]}, "task": {"type": "worker", "index": 0}},
    {"cluster": {"chief": ["a:1", "b:1"], "worker": ["h:1"]}, "task": {"type": "chief", "index": 0}},
    {"cluster": {"gpu": ["h:1"]}, "task": {"type": "worker", "index": 0}},
    {"cluster": {"worker": ["h:1"]}},
])
def test_malformed_configs_raise(bad):
    with pytest.raises(TFConfigError):
        R.from_env(bad if isinstance(bad, str) else json.dumps(bad))


@settings(max_examples=60, deadline=None)
@given(st.text(max_size=40))
def test_garbage_strings_never_crash_unexpectedly(s):
    try:
        R.from_env(s)
    except TFConfigError:
        pass


@settings(max_examples=40, deadline=None)
@given(st.integers(1, 6), st.booleans(), st.data())
def test_ranks_are_a_permutation(nw, with_chief, data):
    workers = [f"h{i}:{1000 + i}" for i in range(nw)]
    chief = "c:999" if with_chief else None
    ranks = []
    if chief:
        ranks.append(R.from_env(make_tf_config(workers, "chief", 0, chief)).id_in_cluster)
    for i in range(nw):
        ranks.append(R.from_env(make_tf_config(workers, "worker", i, chief)).id_in_cluster)
    assert sorted(ranks) == list(range(nw + (1 if chief else 0)))
    chiefs = [R.from_env(make_tf_config(workers, "worker", i, chief)).is_chief for i in range(nw)]
    assert sum(chiefs) == (0 if chief else 1)
