"""Estimator-like driver: RunConfig cadence, checkpoints + auto-resume, summaries, final export, hooks,
model_fn shim."""
import json
import os

import numpy as np
import pytest
import torch

import tfde_b200 as tf
from tfde_b200 import estimator as E
from tfde_b200.ckpt import checkpoint as ckpt
from tfde_b200.obs.summary import read_scalars
from tfde_b200.parallel.strategy import Strategy


@pytest.fixture(autouse=True)
def _no_tf_config(monkeypatch):
    monkeypatch.delenv("TF_CONFIG", raising=False)
    monkeypatch.setenv("TFDE_DEVICE", "cpu")


def _data(n=512):
    (xtr, ytr), (xte, yte) = tf.data.mnist.load_data(n_train=n, n_test=128)
    return (xtr / 255.0, ytr.astype("int").reshape(-1, 1)), (xte / 255.0, yte.astype("int").reshape(-1, 1))


def input_fn(features, labels, batch_size, mode):
    inputs = features if labels is None else (features, labels)
    ds = tf.data.Dataset.from_tensor_slices(inputs)
    if mode == E.ModeKeys.TRAIN:
        ds = ds.shuffle(1000).repeat().batch(batch_size).prefetch(100)
    else:
        ds = ds.batch(batch_size)
    return ds


def serving_input_fn():
    ph = E.placeholder("float32", [None, 28 * 28])
    return E.export.TensorServingInputReceiver(ph, ph)


def _estimator(model_dir, **cfg):
    config = E.RunConfig(model_dir=model_dir, save_summary_steps=5, log_step_count_steps=5, **cfg)
    with Strategy(device="cpu").scope():
        model = tf.models.build_and_compile_model_b(learning_rate=0.05, seed=4)
    return E.model_to_estimator(keras_model=model, model_dir=model_dir, config=config)


def test_train_and_evaluate_local_checkpoints_summaries_export(tmp_path):
    (xtr, ytr), (xte, yte) = _data()
    d = str(tmp_path / "job")
    est = _estimator(d, save_checkpoints_steps=10)
    train_spec = E.TrainSpec(input_fn=lambda: input_fn(xtr, ytr, 32, E.ModeKeys.TRAIN), max_steps=24.5)
    exporter = E.FinalExporter("exporter", serving_input_fn)
    eval_spec = E.EvalSpec(input_fn=lambda: input_fn(xte, yte, 32, E.ModeKeys.EVAL), steps=None, name="mnist-eval",
                           exporters=[exporter], start_delay_secs=0, throttle_secs=0)
    result, _ = E.train_and_evaluate(est, train_spec, eval_spec)
    # max_steps may be a float (468.75 in the reference): train while step < max_steps
    assert est.global_step == 25 and result["global_step"] == 25
    assert 0 <= result["accuracy"] <= 1 and result["loss"] > 0
    files = os.listdir(d)
    assert "checkpoint" in files and any(f.startswith("events.out.tfevents") for f in files)
    steps = sorted(ckpt.step_of(f) for f in files if f.startswith("model.ckpt-"))
    assert steps == [0, 10, 20, 25]
    assert ckpt.latest_checkpoint(d).endswith("model.ckpt-25.npz")
    sc = read_scalars(d)
    assert "loss" in sc and "global_step/sec" in sc
    assert os.path.isdir(os.path.join(d, "eval_mnist-eval"))
    exp_root = os.path.join(d, "export", "exporter")
    stamp = os.listdir(exp_root)
    assert len(stamp) == 1 and stamp[0].isdigit()
    spec = json.load(open(os.path.join(exp_root, stamp[0], "saved_model.json")))
    assert spec["signature"]["serving_default"]["inputs"]["shape"] == [None, 784]
    served = ckpt.load_saved_model(os.path.join(exp_root, stamp[0]), device="cpu")
    out = served.predict(xte[:5].reshape(5, 784).astype(np.float32))
    assert out.shape == (5, 10) and np.allclose(out.sum(1), 1.0, atol=1e-4)   # softmax head


def test_auto_resume_from_latest_checkpoint_and_keep_max(tmp_path):
    (xtr, ytr), _ = _data()
    d = str(tmp_path / "resume")
    est = _estimator(d, save_checkpoints_steps=4, keep_checkpoint_max=3)
    est.train(lambda: input_fn(xtr, ytr, 32, E.ModeKeys.TRAIN), max_steps=12)
    w_after = est._model(E.ModeKeys.TRAIN).named_weights()
    kept = sorted(ckpt.step_of(f) for f in os.listdir(d) if f.startswith("model.ckpt-"))
    assert kept == [4, 8, 12], "keeps only the last 3"
    est2 = _estimator(d, save_checkpoints_steps=4)   # fresh process analogue: new Estimator, same model_dir
    est2.train(lambda: input_fn(xtr, ytr, 32, E.ModeKeys.TRAIN), max_steps=12)
    assert est2.global_step == 12, "already at max_steps: nothing to do"
    w_resumed = est2._model(E.ModeKeys.TRAIN).named_weights()
    for k in w_after:
        np.testing.assert_array_equal(w_after[k], w_resumed[k])
    est2.train(lambda: input_fn(xtr, ytr, 32, E.ModeKeys.TRAIN), steps=3)
    assert est2.global_step == 15 and est2._model(E.ModeKeys.TRAIN).optimizer.iterations == 15


def test_checkpoint_restores_optimizer_slots_and_step(tmp_path):
    with Strategy(device="cpu").scope():
        m = tf.models.build_and_compile_cnn_model(optimizer=tf.keras.optimizers.Adam(0.01), seed=2)
    x = torch.rand(16, 28, 28, 1)
    y = torch.randint(0, 10, (16,))
    for _ in range(3):
        m.train_on_batch(x, y)
    d = str(tmp_path)
    ckpt.save_checkpoint(m, d, step=3)
    m.train_on_batch(x, y)
    expected = m.named_weights()
    with Strategy(device="cpu").scope():
        m2 = tf.models.build_and_compile_cnn_model(optimizer=tf.keras.optimizers.Adam(0.01), seed=77)
        m2.build()
    assert ckpt.restore_checkpoint(m2, ckpt.latest_checkpoint(d)) == 3
    m2.train_on_batch(x, y)
    for k, v in m2.named_weights().items():
        np.testing.assert_allclose(v, expected[k], rtol=1e-6, atol=1e-7, err_msg=k)


def test_profiler_hook_writes_chrome_trace(tmp_path):
    (xtr, ytr), _ = _data(128)
    est = _estimator(str(tmp_path / "p"), save_checkpoints_steps=100)
    hook = E.ProfilerHook(save_steps=2, output_dir=str(tmp_path / "prof"), show_memory=True)
    est.train(lambda: input_fn(xtr, ytr, 16, E.ModeKeys.TRAIN), hooks=[hook], max_steps=5)
    names = sorted(os.listdir(tmp_path / "prof"))
    assert names == ["timeline-0.json", "timeline-2.json", "timeline-4.json"]
    trace = json.load(open(tmp_path / "prof" / names[0]))
    assert trace["traceEvents"][0]["ph"] == "X"
    # per-layer timeline (Model B): at least one span per executed layer in each direction, plus loss and optimizer
    spans = [e["name"] for e in trace["traceEvents"] if e["ph"] == "X"]
    model = est._models[next(iter(est._models))]
    for node in model.exec_plan:
        if getattr(node, "fused_away", False):
            continue
        assert any(f"/{node.name}/forward" in s for s in spans), (node.name, spans)
    for kind in ("Conv2D", "BatchNormalization", "Dense"):
        assert any(s.startswith("host: " + kind) and "/backward" in s for s in spans), kind
    assert any("cross-entropy" in s or "fused head" in s for s in spans) and any("optimizer" in s for s in spans)
    # the step after a sampled one is back on the normal path
    assert model._tracer is None


def test_custom_model_fn_with_manual_global_batch_scaling(tmp_path):
    """tf2_mnist_distributed.py:65-91 re-expressed: reduction=NONE, reduce_sum * (1/BATCH), minimize."""
    BATCH = 32
    (xtr, ytr), (xte, yte) = _data(256)
    xtr4 = xtr.reshape(-1, 28, 28, 1)

    def model_fn(features, labels, mode):
        model = tf.models.build_model_a(seed=8)
        logits = model(features, training=False)
        if mode == E.ModeKeys.PREDICT:
            return E.EstimatorSpec(mode, predictions={"logits": logits})
        optimizer = tf.keras.optimizers.SGD(learning_rate=0.05)
        loss = tf.keras.losses.SparseCategoricalCrossentropy(
            from_logits=True, reduction=tf.keras.losses.Reduction.NONE)(labels, logits)
        loss = E.reduce_sum(loss) * (1. / BATCH)
        if mode == E.ModeKeys.EVAL:
            return E.EstimatorSpec(mode, loss=loss)
        return E.EstimatorSpec(mode=mode, loss=loss,
                               train_op=optimizer.minimize(loss, E.get_or_create_global_step()))

    est = E.Estimator(model_fn=model_fn, model_dir=str(tmp_path / "mf"),
                      config=E.RunConfig(save_checkpoints_steps=100, train_distribute=Strategy(device="cpu")))
    est.train(lambda: input_fn(xtr4, ytr, BATCH, E.ModeKeys.TRAIN), max_steps=6)
    assert est.global_step == 6 and est._train_global_batch == BATCH
    res = est.evaluate(lambda: input_fn(xte.reshape(-1, 28, 28, 1), yte, 64, E.ModeKeys.EVAL))
    assert res["global_step"] == 6 and res["loss"] > 0


def test_session_config_from_env(monkeypatch):
    monkeypatch.setenv("TF_CONFIG", json.dumps({"cluster": {"master": ["a:1"], "worker": ["b:1"], "ps": ["c:1"]},
                                                "task": {"type": "worker", "index": 0}}))
    cfg = E.session_config_from_env()
    assert cfg.device_filters == ["/job:ps", "/job:worker/task:0"]
    monkeypatch.delenv("TF_CONFIG")
    assert E.session_config_from_env() is None
