"""ParameterServerStrategy over its NVLink transport (csrc/kernels/ps_p2p.cu): real processes -- ps tasks holding
their variable shards in GPU memory, workers pushing with reduction atomics / pulling with peer loads through CUDA
IPC mappings. Runs on one GPU too (all tasks share it; nothing in these kernels waits on another process)."""
import json
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "mp_worker_ps_gpu.py")


def _launch(mode, out_dir, nproc, n_ps, **kw):
    from tfde_b200.launch import launch
    os.makedirs(out_dir, exist_ok=True)
    res = launch([WORKER, mode, str(out_dir)], nproc=nproc, n_ps=n_ps, timeout=300, capture=True,
                 extra_env={"TFDE_NO_TENSORBOARD": "1"}, **kw)
    for task, rc, out in res:
        assert rc == 0, f"{task} failed:\n{(out or '')[-3000:]}"


@pytest.mark.timeout(400)
def test_ps_nvlink_push_pull_closed_form(tmp_path):
    n_workers = 3 if torch.cuda.device_count() >= 4 else 2
    _launch("const", tmp_path, nproc=n_workers, n_ps=2, chief=True)
    chief = json.load(open(tmp_path / "const_chief0.json"))
    assert chief["transport"] == "nvlink", chief
    assert chief["max_err"] < 1e-5 * max(1.0, chief["expect_delta"]) + 2e-6, chief
    for i in range(n_workers - 1):
        w = json.load(open(tmp_path / f"const_worker{i}.json"))
        assert w["transport"] == "nvlink" and w["step_seen"] >= 25


@pytest.mark.timeout(400)
def test_ps_estimator_job_on_gpu_with_mirrored_eval(tmp_path):
    _launch("job", tmp_path, nproc=2, n_ps=1, chief=True, env_style="cluster_spec")
    chief = json.load(open(tmp_path / "job_chief0.json"))
    worker = json.load(open(tmp_path / "job_worker0.json"))
    assert chief["transport"] == "nvlink" and worker["transport"] == "nvlink"
    assert chief["global_step"] >= 60 and worker["global_step"] >= 60
    assert chief["eval"] is not None and 0 < chief["eval"]["loss"] < 3.0
    assert os.path.isdir(tmp_path / "model" / "export" / "exporter")


def test_mirrored_eval_over_local_devices_matches_single_device():
    """eval_distribute=MirroredStrategy() (mnist_keras_distributed.py:243): one process, one model replica per local
    device, every batch split between them, sums combined. Two replicas on the same GPU exercise the same path on a
    one-GPU box."""
    import numpy as np
    import tfde_b200 as tf
    n = torch.cuda.device_count()
    devs = ["cuda:0", "cuda:1"] if n >= 2 else ["cuda:0", "cuda:0"]
    (xtr, ytr), (xte, yte) = tf.data.mnist.load_data(n_train=512, n_test=300)
    xte = (xte.reshape(-1, 784) / 255.0).astype(np.float32)
    single = tf.models.build_and_compile_model_b(0.05, seed=5)
    x = (xtr.reshape(-1, 784) / 255.0).astype(np.float32)
    for i in range(4):
        single.train_on_batch(x[i * 128:(i + 1) * 128], ytr[i * 128:(i + 1) * 128].astype(np.int64))
    ref = single.evaluate(xte, yte.astype(np.int64), batch_size=100, return_dict=True)
    st = tf.distribute.MirroredStrategy(devices=devs)
    assert st.num_replicas_in_sync == 1 and len(st.local_devices) == 2
    with st.scope():
        m = tf.models.build_and_compile_model_b(0.05, seed=5)
    m.set_weights(single.get_weights())
    got = m.evaluate(xte, yte.astype(np.int64), batch_size=100, return_dict=True)
    assert len(m._eval_replicas()) == 2
    assert got["accuracy"] == pytest.approx(ref["accuracy"], abs=1e-6)
    assert got["loss"] == pytest.approx(ref["loss"], rel=1e-5)
