"""TF_CONFIG / CLUSTER_SPEC parsing, role -> rank mapping, device filters."""
import json

import pytest

from tfde_b200.cluster import tf_config as tfc


def test_no_tf_config_is_single_local_worker():
    r = tfc.parse_tf_config(None)
    assert not r.distributed and r.is_chief and r.world_size == 1 and r.rank == 0
    assert r.device_filters() is None


def test_two_workers_rank_and_chief():
    cluster = {"worker": ["localhost:12345", "localhost:23456"]}
    r0 = tfc.parse_tf_config(tfc.make_tf_config(cluster, "worker", 0))
    r1 = tfc.parse_tf_config(tfc.make_tf_config(cluster, "worker", 1))
    assert (r0.rank, r0.world_size, r0.is_chief) == (0, 2, True)
    assert (r1.rank, r1.world_size, r1.is_chief) == (1, 2, False)
    assert r0.master_endpoint == ("127.0.0.1", 12345)


def test_chief_ps_evaluator_roles():
    cluster = {"chief": ["h0:1"], "worker": ["h1:1", "h2:1"], "ps": ["p0:2", "p1:2"], "evaluator": ["e:3"]}
    chief = tfc.parse_tf_config(tfc.make_tf_config(cluster, "chief", 0))
    w1 = tfc.parse_tf_config(tfc.make_tf_config(cluster, "worker", 1))
    ps = tfc.parse_tf_config(tfc.make_tf_config(cluster, "ps", 1))
    ev = tfc.parse_tf_config(tfc.make_tf_config(cluster, "evaluator", 0))
    assert chief.is_chief and chief.rank == 0 and chief.world_size == 3
    assert not w1.is_chief and w1.rank == 2
    assert ps.is_ps and ps.rank == -1
    assert ev.is_evaluator and ev.rank == -1
    assert chief.ps_endpoints() == [("p0", 2), ("p1", 2)]


def test_master_is_chief_legacy():
    cluster = {"master": ["m:1"], "worker": ["w:1"], "ps": ["p:1"]}
    m = tfc.parse_tf_config(tfc.make_tf_config(cluster, "master", 0))
    w = tfc.parse_tf_config(tfc.make_tf_config(cluster, "worker", 0))
    assert m.is_chief and not w.is_chief and m.rank == 0 and w.rank == 1


def test_device_filters_match_reference_rules():
    cluster = {"master": ["m:1"], "worker": ["w0:1", "w1:1"], "ps": ["p:1"]}
    f = lambda t, i: tfc.device_filters_from_tf_config(json.loads(tfc.make_tf_config(cluster, t, i)))
    assert f("master", 0) == ["/job:ps", "/job:master"]
    assert f("worker", 1) == ["/job:ps", "/job:worker/task:1"]
    assert f("ps", 0) is None
    assert tfc.device_filters_from_tf_config({}) is None


def test_cluster_spec_shim_sets_tf_config_and_prints(capsys):
    env = {"CLUSTER_SPEC": json.dumps({"worker": ["a:1", "b:1"], "ps": ["c:1"]}), "JOB_NAME": "worker",
           "TASK_INDEX": "1"}
    out = tfc.synthesize_tf_config_from_cluster_spec(env)
    cfg = json.loads(env["TF_CONFIG"])
    assert cfg["task"] == {"type": "worker", "index": 1} and cfg["cluster"]["ps"] == ["c:1"]
    assert out == env["TF_CONFIG"]
    assert "Distribution enabled" in capsys.readouterr().out
    env2 = {}
    assert tfc.synthesize_tf_config_from_cluster_spec(env2) is None and "TF_CONFIG" not in env2
    assert "Distribution is not enabled" in capsys.readouterr().out


@pytest.mark.parametrize("bad", ["{not json", json.dumps({"cluster": {"wrkr": ["a:1"]}}),
                                 json.dumps({"cluster": {"worker": ["a:1"]}, "task": {"type": "worker", "index": 3}}),
                                 json.dumps({"cluster": {"worker": ["a:1"]}, "task": {"type": "ps", "index": 0}})])
def test_invalid_tf_config_raises(bad):
    with pytest.raises(tfc.ClusterConfigError):
        tfc.parse_tf_config(bad)


def test_torchrun_env_synthesis():
    s = tfc.tf_config_from_torchrun_env({"RANK": "2", "WORLD_SIZE": "4", "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": "29500"})
    r = tfc.parse_tf_config(s)
    assert r.rank == 2 and r.world_size == 4
