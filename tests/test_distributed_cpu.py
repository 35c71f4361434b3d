"""Distributed plumbing on CPU/gloo with real processes (SURVEY.md section 4 tier 3)."""
import json
import os
import sys

import numpy as np
import pytest
import torch

import tfde_b200 as tf
from tfde_b200.launch import launch
from tfde_b200.parallel.strategy import Strategy

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "mp_worker_cpu.py")


def _run(mode, out_dir, **kw):
    res = launch([WORKER, mode, str(out_dir)], extra_env={"TFDE_DEVICE": "cpu", "OMP_NUM_THREADS": "2"},
                 timeout=240, capture=True, **kw)
    for task, rc, out in res:
        assert rc == 0, f"{task} failed:\n{out[-3000:]}"
    return res


@pytest.mark.timeout(300)
def test_two_workers_tf_config_sync_training_matches_single_process(tmp_path):
    _run("mwms", tmp_path, nproc=2)
    r0 = json.load(open(tmp_path / "result_rank0.json"))
    r1 = json.load(open(tmp_path / "result_rank1.json"))
    assert r0["world"] == r1["world"] == 2 and r0["chief"] and not r1["chief"]
    # identical replicas: same weights after N steps, same aggregated metrics on every worker
    w0, w1 = np.load(tmp_path / "weights_rank0.npz"), np.load(tmp_path / "weights_rank1.npz")
    for k in w0.files:
        np.testing.assert_array_equal(w0[k], w1[k], err_msg=k)
    assert r0["loss"] == r1["loss"] and r0["eval"] == r1["eval"]
    # chief-only side effect
    assert os.path.exists(tmp_path / "chief_weights.npz")
    # equals one process training on the global batch (seed of the chief = 100)
    (xtr, ytr), _ = tf.data.mnist.load_data(n_train=512, n_test=128)
    with Strategy(device="cpu").scope():
        m = tf.models.build_and_compile_cnn_model(learning_rate=0.05, seed=100)
    ds = tf.data.Dataset.from_tensor_slices((xtr[..., None], ytr.astype(np.int64))).batch(32)
    h = m.fit(ds, epochs=2, steps_per_epoch=4, verbose=0)
    np.testing.assert_allclose(r0["loss"], h.history["loss"], rtol=1e-4)
    for k, v in m.named_weights().items():
        np.testing.assert_allclose(w0[k], v, rtol=1e-3, atol=1e-5, err_msg=k)


@pytest.mark.timeout(300)
def test_parameter_server_job_with_cluster_spec_env(tmp_path):
    """1 chief + 1 worker + 1 ps started from CLUSTER_SPEC/JOB_NAME/TASK_INDEX (the reference's YARN-style
    contract): async updates, global_step counted on the ps, ps exits when training ends, chief exports."""
    _run("ps", tmp_path, nproc=2, n_ps=1, chief=True, env_style="cluster_spec")
    chief = json.load(open(tmp_path / "result_chief0.json"))
    worker = json.load(open(tmp_path / "result_worker0.json"))
    assert chief["global_step"] >= 30 and worker["global_step"] >= 30
    assert worker["filters"] == ["/job:ps", "/job:worker/task:0"]
    assert chief["eval"] is not None and chief["eval"]["loss"] > 0
    model_dir = tmp_path / "model"
    assert any(f.startswith("model.ckpt-") for f in os.listdir(model_dir))
    assert os.path.isdir(model_dir / "export" / "exporter")
