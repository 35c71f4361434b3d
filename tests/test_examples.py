"""The three reference scripts, re-expressed on this framework, run end to end (CPU, short)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, env_extra, timeout=240):
    env = dict(os.environ, TFDE_DEVICE="cpu", OMP_NUM_THREADS="4", TFDE_NO_TENSORBOARD="1", **env_extra)
    env.pop("TF_CONFIG", None)
    r = subprocess.run([sys.executable] + args, cwd=ROOT, env=env, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    return r.stdout + r.stderr


@pytest.mark.timeout(300)
def test_distributed_with_keras_example_single_worker():
    out = _run(["examples/distributed_with_keras.py"], {})
    assert "Epoch 3/3" in out and "5/5" in out


@pytest.mark.timeout(300)
def test_mnist_keras_distributed_example_local(tmp_path):
    out = _run(["examples/mnist_keras_distributed.py", "--working-dir", str(tmp_path / "wd"), "--batch-size", "64",
                "--max-steps", "12", "--unknown-flag", "ignored"], {})
    assert "Distribution is not enabled" in out and "Total params: 250,950" in out
    assert os.path.exists(tmp_path / "wd" / "checkpoint")
    assert os.path.isdir(tmp_path / "wd" / "export" / "exporter")


@pytest.mark.timeout(300)
def test_tf2_mnist_distributed_example_both_models(tmp_path):
    _run(["examples/tf2_mnist_distributed.py"], {"TFDE_MODEL_DIR": str(tmp_path / "b"), "TFDE_MAX_STEPS": "6"})
    assert os.path.exists(tmp_path / "b" / "checkpoint")
    _run(["examples/tf2_mnist_distributed.py", "--model-fn"],
         {"TFDE_MODEL_DIR": str(tmp_path / "a"), "TFDE_MAX_STEPS": "6"})
    assert os.path.exists(tmp_path / "a" / "checkpoint")


@pytest.mark.timeout(420)
def test_mnist_keras_distributed_example_ps_cluster(tmp_path):
    """The reference's own deployment shape (mnist_keras_distributed.py:221-243): CLUSTER_SPEC / JOB_NAME /
    TASK_INDEX with ps, master and worker tasks, ParameterServerStrategy for training and MirroredStrategy() as
    eval_distribute constructed in EVERY task. The ps task must serve (not die in MirroredStrategy), the training
    tasks must not wait in a synchronous rendezvous."""
    sys.path.insert(0, ROOT)
    from tfde_b200 import launch
    wd = str(tmp_path / "wd")
    env = {"TFDE_DEVICE": "cpu", "OMP_NUM_THREADS": "2", "TFDE_NO_TENSORBOARD": "1", "TFDE_RENDEZVOUS_TIMEOUT": "60"}
    old_tf = os.environ.pop("TF_CONFIG", None)
    try:
        res = launch.launch([os.path.join(ROOT, "examples", "mnist_keras_distributed.py"), "--working-dir", wd,
                             "--batch-size", "64", "--max-steps", "16"], nproc=2, n_ps=1, chief=False,
                            env_style="cluster_spec", extra_env=env, timeout=360, capture=True)
    finally:
        if old_tf is not None:
            os.environ["TF_CONFIG"] = old_tf
    for task, rc, out in res:
        assert rc == 0, f"{task}: rc={rc}\n{(out or '')[-3000:]}"
    assert any("Distribution enabled" in (out or "") for _, _, out in res)
    assert os.path.exists(os.path.join(wd, "checkpoint"))
