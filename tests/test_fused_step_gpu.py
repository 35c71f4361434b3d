"""The whole-step kernel (csrc/kernels/fused_step.cu: forward + loss + backward + optimizer of the reference's Keras
MNIST CNN in one cooperative launch) against two independent oracles: the PyTorch CPU restatement of the model
(ops/reference.py, driven layer by layer) and the layer-by-layer GPU kernels of round 1."""
import numpy as np
import pytest
import torch

import tfde_b200 as tf
from tfde_b200.parallel.strategy import Strategy

pytestmark = pytest.mark.gpu


def _data(n, seed=0, u8=True):
    g = torch.Generator().manual_seed(seed)
    x = torch.randint(0, 256, (n, 28, 28, 1), generator=g, dtype=torch.uint8)
    y = torch.randint(0, 10, (n,), generator=g)
    return (x if u8 else x.float() / 255.0), y


def _opt(kind):
    return {"sgd": lambda: tf.keras.optimizers.SGD(0.05),
            "momentum": lambda: tf.keras.optimizers.SGD(0.05, momentum=0.9),
            "nesterov": lambda: tf.keras.optimizers.SGD(0.05, momentum=0.9, nesterov=True),
            "adam": lambda: tf.keras.optimizers.Adam(0.002)}[kind]()


def _close(got, want, tol=2e-3):
    """Weights after several optimizer steps: norm-wise per tensor. (Element-wise tolerances are meaningless for the
    few weights whose gradient is ~0: one ReLU / arg-max decision flipping under a different summation order, or
    Adam's g/sqrt(v) on a noise-level gradient, moves them by a full learning-rate step.)"""
    for a, b in zip(got, want):
        err = float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-12))
        assert err < tol, err


def _train(device, kind, steps, x, y, bs, seed=7):
    with Strategy(device=device).scope():
        m = tf.models.build_and_compile_cnn_model(optimizer=_opt(kind), seed=seed)
    losses = []
    for i in range(steps):
        lo = (i * bs) % (len(x) - bs + 1)
        m.train_on_batch(x[lo:lo + bs], y[lo:lo + bs])
        losses.append(m.last_step_loss())
    return m, losses


@pytest.mark.parametrize("kind", ["sgd", "momentum", "nesterov", "adam"])
def test_whole_step_matches_cpu_reference(kind):
    x, y = _data(512)
    mc, lc = _train("cpu", kind, 8, x, y, 64)
    mg, lg = _train("cuda", kind, 8, x, y, 64)
    assert mg._fused is not None, "the whole-step kernel must be the path that ran"
    np.testing.assert_allclose(lg, lc, rtol=2e-4, atol=2e-5)
    if kind == "adam":
        # Adam turns the first noise-level gradient of a so-far-dead unit into a full learning-rate step, so a ReLU
        # decision that flips under a different summation order moves a handful of weights by ~lr: bound the size of
        # any deviation by one such step and the number of deviating weights
        for a, b in zip(mg.get_weights(), mc.get_weights()):
            d = np.abs(a - b)
            assert d.max() < 1.5 * 0.002, d.max()
            assert (d > 2e-4).mean() < 2e-2, (d > 2e-4).mean()
    else:
        _close(mg.get_weights(), mc.get_weights())


@pytest.mark.parametrize("bs", [1, 7, 37, 64])
def test_whole_step_batch_sizes_and_float_input(bs):
    x, y = _data(256, seed=3, u8=False)
    mc, lc = _train("cpu", "sgd", 4, x, y, bs)
    mg, lg = _train("cuda", "sgd", 4, x, y, bs)
    assert mg._fused is not None
    np.testing.assert_allclose(lg, lc, rtol=2e-4, atol=2e-5)
    _close(mg.get_weights(), mc.get_weights())


def test_whole_step_matches_layerwise_kernels(monkeypatch):
    x, y = _data(512, seed=5)
    mf, lf = _train("cuda", "sgd", 10, x, y, 64)
    monkeypatch.setenv("TFDE_FUSED_STEP", "0")
    ml, ll = _train("cuda", "sgd", 10, x, y, 64)
    assert mf._fused is not None and ml._fused is None
    np.testing.assert_allclose(lf, ll, rtol=1e-4, atol=1e-5)
    _close(mf.get_weights(), ml.get_weights(), 1e-3)
    # gradients of the last step are left in the gradient arena by both paths
    _close([mf.arena.g.cpu().numpy()], [ml.arena.g.cpu().numpy()], 1e-3)


def test_steps_per_execution_equals_single_steps():
    """K steps in one launch (each with its own batch) == K launches of one step, bit for bit."""
    x, y = _data(64 * 6, seed=9)
    m1, l1 = _train("cuda", "adam", 6, x, y, 64)
    with Strategy(device="cuda").scope():
        m2 = tf.models.build_and_compile_cnn_model(optimizer=_opt("adam"), seed=7)
    m2.train_on_batches(x[:64 * 2], y[:64 * 2], steps=2)
    la = m2.last_losses()
    m2.train_on_batches(x[64 * 2:].reshape(4, 64, 28, 28, 1), y[64 * 2:].reshape(4, 64))
    lb = m2.last_losses()
    assert m2.optimizer.iterations == 6
    assert la + lb == l1
    for a, b in zip(m1.get_weights(), m2.get_weights()):
        assert np.array_equal(a, b)
    # device-resident inputs take the same kernel
    with Strategy(device="cuda").scope():
        m3 = tf.models.build_and_compile_cnn_model(optimizer=_opt("adam"), seed=7)
    m3.train_on_batches(x.cuda(), y.cuda(), steps=6)
    for a, b in zip(m1.get_weights(), m3.get_weights()):
        assert np.array_equal(a, b)


def test_fit_with_steps_per_execution_learns():
    (xtr, ytr), (xte, yte) = tf.data.mnist.load_data(n_train=2048, n_test=512)
    m = tf.models.build_and_compile_cnn_model(learning_rate=0.05, seed=1)
    m.compile(optimizer=tf.keras.optimizers.SGD(0.05),
              loss=tf.keras.losses.SparseCategoricalCrossentropy(from_logits=True), metrics=["accuracy"],
              steps_per_execution=8)
    ds = tf.data.Dataset.from_tensor_slices((xtr[..., None], ytr.astype(np.int64))).shuffle(2048, seed=1).repeat().batch(64)
    h = m.fit(ds, epochs=3, steps_per_epoch=36, verbose=0)     # 36 = 4 full groups of 8 + one of 4
    assert m._fused is not None
    assert m.optimizer.iterations == 108
    assert h.history["loss"][-1] < h.history["loss"][0]
    loss, acc = m.evaluate(xte[..., None], yte.astype(np.int64), batch_size=128)
    assert acc > 0.9, (loss, acc)


def test_out_of_range_label_poisons_the_loss_not_the_weights():
    x, y = _data(64, seed=11)
    y = y.clone()
    y[5] = 12
    with Strategy(device="cuda").scope():
        m = tf.models.build_and_compile_cnn_model(learning_rate=0.05, seed=2)
    m.train_on_batch(x, y)
    assert np.isnan(m.last_step_loss())
    for w in m.get_weights():
        assert np.isfinite(w).all()


def test_metrics_and_learning_rate_change_between_launches():
    x, y = _data(256, seed=13)
    with Strategy(device="cuda").scope():
        m = tf.models.build_and_compile_cnn_model(learning_rate=0.05, seed=4)
    m.train_on_batches(x, y, steps=4)
    met = m.read_metrics(4)
    ls = m.last_losses()
    assert met["loss"] == pytest.approx(sum(ls) / 4, rel=1e-5)
    assert 0.0 <= met["accuracy"] <= 1.0
    w0 = [w.copy() for w in m.get_weights()]
    m.optimizer.set_learning_rate(0.0)
    m.train_on_batches(x, y, steps=4)
    for a, b in zip(w0, m.get_weights()):
        assert np.array_equal(a, b), "lr = 0 must leave the weights untouched"


@pytest.mark.parametrize("kind,bs", [("sgd", 64), ("momentum", 37)])
def test_whole_step_tensor_core_variant(monkeypatch, kind, bs):
    """TFDE_FUSED_TC=1 (experimental, opt-in): the three per-CTA GEMMs as TF32 tcgen05.mma with the accumulators in
    TMEM. TF32 keeps 10 mantissa bits of each operand: the loss follows the fp32 CPU restatement to ~1e-4, but hidden
    units whose pre-activation is within TF32 error of zero flip their ReLU, so single weights move by a whole
    one-sample gradient; the check is therefore on the loss trajectory and on the bulk of the weights."""
    monkeypatch.setenv("TFDE_FUSED_TC", "1")
    monkeypatch.setenv("TFDE_FUSED_TC_FORCE", "1")
    x, y = _data(512, seed=21)
    mc, lc = _train("cpu", kind, 8, x, y, bs)
    mg, lg = _train("cuda", kind, 8, x, y, bs)
    assert mg._fused is not None and mg._fused.tensor_core
    mg._fused.check()
    np.testing.assert_allclose(lg, lc, rtol=2e-3, atol=2e-4)
    for a, b in zip(mg.get_weights(), mc.get_weights()):
        d = np.abs(a - b) / max(np.abs(b).max(), 1e-12)
        assert np.median(d) < 3e-2 and d.max() < 0.2, (np.median(d), d.max())     # TF32: sums with heavy cancellation
    # multi-step launch == single steps, bit for bit, in this variant too
    with Strategy(device="cuda").scope():
        m2 = tf.models.build_and_compile_cnn_model(optimizer=_opt(kind), seed=7)
    m2.train_on_batches(x[:8 * bs], y[:8 * bs], steps=8)
    for a, b in zip(mg.get_weights(), m2.get_weights()):
        assert np.array_equal(a, b)


def test_epoch_counters_wrap_around_2_pow_32():
    """The kernel's barrier / ticket / epoch counters only ever grow (they make the kernel replayable without host
    resets) and are compared modulo 2^32. Start them 20 steps before the wrap and train across it: the result must be
    bit-identical to a run that starts at zero."""
    x, y = _data(64 * 8, seed=31)
    ref_m, ref_l = _train("cuda", "momentum", 8, x, y, 64)          # counters start at 0
    with Strategy(device="cuda").scope():
        m = tf.models.build_and_compile_cnn_model(optimizer=_opt("momentum"), seed=7)
    m.train_on_batch(x[:64], y[:64])                                 # builds the plan (epoch 1)
    m.sync()
    plan = m.fused_plan()
    assert plan is not None
    # rebuild the same model so the weights are fresh, then plant the counters: every word is a multiple of the epoch
    with Strategy(device="cuda").scope():
        m = tf.models.build_and_compile_cnn_model(optimizer=_opt("momentum"), seed=7)
    m.train_on_batches(x[:64].cuda(), y[:64].cuda(), steps=1)       # epoch 1 on the new plan (its own state words)
    m.sync()
    plan = m.fused_plan()
    E0 = (1 << 32) - 3                                               # 3 steps before the wrap
    grid = 169
    st = torch.zeros_like(plan.state, dtype=torch.int64)
    st[0] = (E0 * 2 * grid) % (1 << 32)                              # ST_BAR
    st[32] = (E0 * 13) % (1 << 32)                                   # ST_TICKET
    st[64] = E0                                                      # ST_DONE
    st[96] = E0                                                      # ST_EPOCH0
    for g in range(13):
        st[128 + 32 * g] = (E0 * 13) % (1 << 32)                     # ST_GROUP
    plan.state.copy_(torch.where(st >= (1 << 31), st - (1 << 32), st).to(torch.int32))
    m.train_on_batches(x[64:64 * 5].cuda(), y[64:64 * 5].cuda(), steps=4)   # crosses 2^32 inside one launch
    for k in range(5, 8):
        m.train_on_batch(x[64 * k:64 * (k + 1)], y[64 * k:64 * (k + 1)])
    m.sync()
    plan.check()
    assert int(plan.state[96].item()) & 0xFFFFFFFF == (E0 + 7) % (1 << 32)
    for a, b in zip(m.get_weights(), ref_m.get_weights()):
        assert np.array_equal(a, b)


def test_streamed_inputs_equal_staged_inputs(monkeypatch):
    """End-to-end executions with the kernel launched ahead of its host->device copies (it takes up every step when that
    step's ready tag has landed) == all copies first, then the launch: bit for bit, over many back-to-back executions
    that re-use the two staging slots (growing tags), from pinned and from pageable host memory."""
    x, y = _data(64 * 40, seed=11)
    xp, yp = x.pin_memory(), y.pin_memory()
    res = {}
    for mode in ("1", "0"):
        monkeypatch.setenv("TFDE_FUSED_STREAM", mode)
        with Strategy(device="cuda").scope():
            m = tf.models.build_and_compile_cnn_model(optimizer=_opt("momentum"), seed=7)
        losses = []
        for rep in range(3):                                   # 3 x (8 + 8 + 4) steps from pinned memory, pipelined
            for lo, k in ((0, 8), (8, 8), (16, 4)):
                m.train_on_batches(xp[lo * 64:(lo + k) * 64], yp[lo * 64:(lo + k) * 64], steps=k)
                losses += m.last_losses()
        m.train_on_batches(x[20 * 64:], y[20 * 64:], steps=20)  # pageable source, one long execution
        losses += m.last_losses()
        m.fused_plan().check()
        res[mode] = (losses, m.get_weights())
    assert len(res["1"][0]) == 80 and res["1"][0] == res["0"][0]
    assert all(np.isfinite(v) for v in res["1"][0])
    for a, b in zip(res["1"][1], res["0"][1]):
        assert np.array_equal(a, b)
