"""End-to-end on one GPU: the hand-written kernel path must follow the PyTorch-CPU path step by step
(same init, same data), eager and CUDA-graph replay must agree bit for bit, and the loss must go down."""
import numpy as np
import pytest
import torch

import tfde_b200 as tf
from tfde_b200.parallel.strategy import Strategy

pytestmark = pytest.mark.gpu


def _data(n, flat=False, seed=0):
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(n, 784 if flat else 28 * 28, generator=g).reshape((n, 784) if flat else (n, 28, 28, 1))
    y = torch.randint(0, 10, (n,), generator=g)
    return x, y


def _train(builder, device, steps, x, y, bs, graph=True):
    with Strategy(device=device).scope():
        m = builder()
    m.use_cuda_graph = graph
    losses = []
    for i in range(steps):
        lo = (i * bs) % (len(x) - bs + 1)
        m.train_on_batch(x[lo:lo + bs], y[lo:lo + bs])
        losses.append(m.last_step_loss())
    return m, losses


@pytest.mark.parametrize("opt", ["sgd", "adam"])
def test_model_a_matches_cpu_reference(opt):
    mk_opt = (lambda: tf.keras.optimizers.SGD(0.05)) if opt == "sgd" else (lambda: tf.keras.optimizers.Adam(0.002))
    builder = lambda: tf.models.build_and_compile_cnn_model(optimizer=mk_opt(), seed=7)
    x, y = _data(512)
    mc, lc = _train(builder, "cpu", 8, x, y, 64)
    mg, lg = _train(builder, "cuda", 8, x, y, 64)
    np.testing.assert_allclose(lg, lc, rtol=2e-3, atol=2e-4)
    for a, b in zip(mg.get_weights(), mc.get_weights()):
        np.testing.assert_allclose(a, b, rtol=5e-3, atol=2e-4)


def test_model_a_graph_equals_eager(monkeypatch):
    monkeypatch.setenv("TFDE_FUSED_STEP", "0")     # the layer-by-layer path (the whole-step kernel needs no graph)
    builder = lambda: tf.models.build_and_compile_cnn_model(learning_rate=0.05, seed=3)
    x, y = _data(256)
    m1, l1 = _train(builder, "cuda", 10, x, y, 64, graph=True)
    m2, l2 = _train(builder, "cuda", 10, x, y, 64, graph=False)
    assert m1.graph_captured, "graph was not captured"
    assert l1 == l2, "CUDA-graph replay and eager launches must be bit-identical"
    for a, b in zip(m1.get_weights(), m2.get_weights()):
        assert np.array_equal(a, b)


def test_model_a_uint8_input_fused_scale():
    builder = lambda: tf.models.build_and_compile_cnn_model(learning_rate=0.05, seed=5)
    g = torch.Generator().manual_seed(1)
    xu = torch.randint(0, 256, (128, 28, 28, 1), generator=g, dtype=torch.uint8)
    y = torch.randint(0, 10, (128,), generator=g)
    _, lu = _train(builder, "cuda", 4, xu, y, 64)
    _, lf = _train(builder, "cuda", 4, xu.float() / 255.0, y, 64)
    np.testing.assert_allclose(lu, lf, rtol=1e-5, atol=1e-6)


def test_model_b_matches_cpu_reference():
    builder = lambda: tf.models.build_and_compile_model_b(learning_rate=0.05, seed=11)
    x, y = _data(512, flat=True)
    mc, lc = _train(builder, "cpu", 6, x, y.reshape(-1, 1), 128)
    mg, lg = _train(builder, "cuda", 6, x, y.reshape(-1, 1), 128)
    np.testing.assert_allclose(lg, lc, rtol=5e-3, atol=5e-4)
    wc, wg = mc.named_weights(), mg.named_weights()
    for k in wc:
        np.testing.assert_allclose(wg[k], wc[k], rtol=2e-2, atol=5e-3, err_msg=k)
    # inference path (moving statistics, no dropout)
    ec = mc.evaluate(x[:256], y[:256].reshape(-1, 1), batch_size=128, return_dict=True)
    eg = mg.evaluate(x[:256], y[:256].reshape(-1, 1), batch_size=128, return_dict=True)
    assert eg["loss"] == pytest.approx(ec["loss"], rel=5e-3)
    assert eg["accuracy"] == pytest.approx(ec["accuracy"], abs=0.02)


def test_fit_and_evaluate_on_gpu_learns():
    (xtr, ytr), (xte, yte) = tf.data.mnist.load_data(n_train=2048, n_test=512)
    m = tf.models.build_and_compile_cnn_model(learning_rate=0.05, seed=1)
    assert m.device.type == "cuda"
    ds = tf.data.Dataset.from_tensor_slices((xtr[..., None], ytr.astype(np.int64))).shuffle(2048, seed=1).repeat().batch(64)
    h = m.fit(ds, epochs=3, steps_per_epoch=32, verbose=0)
    assert h.history["loss"][-1] < h.history["loss"][0]
    loss, acc = m.evaluate(xte[..., None], yte.astype(np.int64), batch_size=128)
    assert acc > 0.9, (loss, acc)


def test_graft_smoke():
    import __graft_entry__ as ge
    ge.smoke()


def test_tiny_resnet_matches_cpu_reference():
    """Residual blocks, BN with gamma, 7x7/s2 + 3x3/s2 'same' max-pool, GAP, Dense head on the GPU kernels."""
    def builder():
        m = tf.models.build_resnet((1, 2), width=8, classes=10, input_shape=(32, 32, 3), seed=13)
        m.compile(optimizer=tf.keras.optimizers.SGD(0.05, momentum=0.9),
                  loss=tf.keras.losses.SparseCategoricalCrossentropy(from_logits=True), metrics=["accuracy"])
        return m
    g = torch.Generator().manual_seed(2)
    x = torch.rand(64, 32, 32, 3, generator=g)
    y = torch.randint(0, 10, (64,), generator=g)
    mc, lc = _train(builder, "cpu", 4, x, y, 16)
    mg, lg = _train(builder, "cuda", 4, x, y, 16)
    np.testing.assert_allclose(lg, lc, rtol=1e-2, atol=1e-3)


def test_profiler_hook_per_kernel_timeline_on_gpu(tmp_path):
    """ProfilerHook (mnist_keras_distributed.py:235-237): the sampled step's Chrome trace has a device span per
    layer and direction that names the native entry points it launched; the whole-step kernel of Model A is
    resolved into its in-kernel phases."""
    import json
    from tfde_b200 import estimator as E
    x, y = _data(128, flat=True)
    hook = E.ProfilerHook(save_steps=3, output_dir=str(tmp_path / "b"))
    mb = tf.models.build_and_compile_model_b(learning_rate=0.05, seed=1)
    hook.begin()
    for step in range(4):
        hook.before_run(step, mb)
        mb.train_on_batch(x, y.reshape(-1, 1))
        hook.after_run(step, mb)
    ev = json.load(open(tmp_path / "b" / "timeline-3.json"))["traceEvents"]
    dev = [e for e in ev if e.get("pid") == 1 and e["ph"] == "X" and e.get("tid") == 0]
    names = [e["name"] for e in dev]
    assert sum("/forward" in n for n in names) >= 8 and sum("/backward" in n for n in names) >= 8, names
    kernels = {k for e in dev for k in e.get("args", {}).get("kernels", [])}
    assert any(k.startswith(("tfde_conv2d", "tfde_smallconv")) for k in kernels) and any(k.startswith("tfde_bn") for k in kernels), kernels
    assert all(e["dur"] >= 0 for e in dev)
    # Model A: one kernel per step, its phases come from the in-kernel stamps
    hook = E.ProfilerHook(save_steps=2, output_dir=str(tmp_path / "a"))
    ma = tf.models.build_and_compile_cnn_model(learning_rate=0.05, seed=1)
    xa, ya = _data(64)
    hook.begin()
    for step in range(3):
        hook.before_run(step, ma)
        ma.train_on_batch(xa, ya)
        hook.after_run(step, ma)
    ev = json.load(open(tmp_path / "a" / "timeline-2.json"))["traceEvents"]
    phases = [e["name"] for e in ev if e.get("tid") == 1]
    assert any("P1 conv" in n for n in phases) and any("two-level ticket" in n for n in phases), phases
