"""CPU path: explicit forward/backward of every layer must equal PyTorch autograd on the same math; Keras
defaults (initialisers, parameter counts, naming, layouts) must match the reference's models."""
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F

import tfde_b200 as tf
from tfde_b200.ops import reference as ref
from tfde_b200.parallel.strategy import Strategy


def cpu_scope():
    return Strategy(device="cpu").scope()


def test_model_a_summary_counts_and_names():
    with cpu_scope():
        m = tf.models.build_model_a(seed=0)
        m.build()
    assert m.count_params() == 347_146
    names = [v.name for v in m.variables]
    assert names == ["conv2d/kernel", "conv2d/bias", "dense/kernel", "dense/bias", "dense_1/kernel", "dense_1/bias"]
    shapes = [v.shape for v in m.variables]
    assert shapes == [(3, 3, 1, 32), (32,), (5408, 64), (64,), (64, 10), (10,)]
    text = m.summary(print_fn=lambda s: None)
    assert "Total params: 347,146" in text


def test_model_b_summary_counts_and_names():
    with cpu_scope():
        m = tf.models.build_model_b(seed=0)
        m.build()
    train = sum(v.size for v in m.trainable_variables)
    frozen = sum(v.size for v in m.non_trainable_variables)
    assert (train, frozen) == (250_466, 484)
    names = [v.name for v in m.variables]
    assert names[:4] == ["conv2d/kernel", "batch_normalization/beta", "batch_normalization/moving_mean",
                         "batch_normalization/moving_variance"]
    assert "batch_normalization_3/beta" in names and "dense_1/bias" in names
    assert not any("gamma" in n for n in names), "scale=False -> no gamma"


def test_glorot_uniform_limits_and_zero_bias():
    with cpu_scope():
        m = tf.models.build_model_a(seed=3)
        m.build()
    w = m.named_weights()
    lim_conv = math.sqrt(6.0 / (9 * 1 + 9 * 32))
    lim_dense = math.sqrt(6.0 / (5408 + 64))
    assert np.abs(w["conv2d/kernel"]).max() <= lim_conv + 1e-7
    assert np.abs(w["dense/kernel"]).max() <= lim_dense + 1e-7
    assert np.abs(w["dense/kernel"]).max() > 0.9 * lim_dense
    assert not w["conv2d/bias"].any() and not w["dense_1/bias"].any()


def test_same_padding_rule():
    assert ref.same_padding(28, 3, 1) == (28, 1, 1)
    assert ref.same_padding(28, 6, 2) == (14, 2, 2)
    assert ref.same_padding(14, 6, 2) == (7, 2, 2)
    assert ref.same_padding(7, 3, 2) == (4, 1, 1)
    assert ref.same_padding(8, 3, 2) == (4, 0, 1)   # extra pixel goes to the high side


def _autograd_model_a(w, x, y):
    """Model A in plain torch with autograd, weights in Keras layout."""
    ws = {k: torch.tensor(v, dtype=torch.float64, requires_grad=True) for k, v in w.items()}
    h = F.conv2d(x.permute(0, 3, 1, 2), ws["conv2d/kernel"].permute(3, 2, 0, 1), ws["conv2d/bias"])
    h = F.max_pool2d(torch.relu(h), 2).permute(0, 2, 3, 1).reshape(x.shape[0], -1)
    h = torch.relu(h @ ws["dense/kernel"] + ws["dense/bias"])
    z = h @ ws["dense_1/kernel"] + ws["dense_1/bias"]
    loss = F.cross_entropy(z, y, reduction="sum") / x.shape[0]
    loss.backward()
    return loss.item(), {k: v.grad for k, v in ws.items()}


def test_model_a_gradients_match_autograd():
    with cpu_scope():
        m = tf.models.build_and_compile_cnn_model(learning_rate=0.0, seed=5)
    g = torch.Generator().manual_seed(0)
    x = torch.rand(16, 28, 28, 1, generator=g)
    y = torch.randint(0, 10, (16,), generator=g)
    w0 = m.named_weights()
    m.train_on_batch(x, y)
    loss, grads = _autograd_model_a(w0, x.double(), y)
    assert m.last_step_loss() == pytest.approx(loss, rel=1e-5)
    for v in m.trainable_variables:
        np.testing.assert_allclose(v.grad.numpy(), grads[v.name].numpy(), rtol=2e-4, atol=1e-6, err_msg=v.name)


def test_model_b_gradients_match_autograd():
    """BN (batch stats) + fused ReLU + softmax/probability loss; dropout disabled via rate=0 clone."""
    l = tf.keras.layers
    with cpu_scope():
        m = tf.keras.Sequential([
            l.Reshape(input_shape=(784,), target_shape=(28, 28, 1)),
            l.Conv2D(6, 3, padding="same", use_bias=False), l.BatchNormalization(scale=False), l.Activation("relu"),
            l.Conv2D(12, 6, padding="same", use_bias=False, strides=2), l.BatchNormalization(scale=False), l.Activation("relu"),
            l.Flatten(), l.Dense(20, use_bias=False), l.BatchNormalization(scale=False), l.Activation("relu"),
            l.Dropout(0.0), l.Dense(10, activation="softmax")], seed=9)
        m.compile(optimizer=tf.keras.optimizers.SGD(0.0), loss="sparse_categorical_crossentropy", metrics=["accuracy"])
    g = torch.Generator().manual_seed(1)
    x = torch.rand(8, 784, generator=g)
    y = torch.randint(0, 10, (8, 1), generator=g)
    w0 = {k: torch.tensor(v, dtype=torch.float64, requires_grad=True) for k, v in m.named_weights().items()}
    m.train_on_batch(x, y)

    def bn(h, beta):
        dims = tuple(range(h.dim() - 1))
        mu, var = h.mean(dims), h.var(dims, unbiased=False)
        return (h - mu) / torch.sqrt(var + 1e-3) + beta

    def conv(h, k, stride, pads):
        hp = F.pad(h.permute(0, 3, 1, 2), (pads[2], pads[3], pads[0], pads[1]))
        return F.conv2d(hp, k.permute(3, 2, 0, 1), stride=stride).permute(0, 2, 3, 1)

    h = x.double().reshape(8, 28, 28, 1)
    h = torch.relu(bn(conv(h, w0["conv2d/kernel"], 1, (1, 1, 1, 1)), w0["batch_normalization/beta"]))
    h = torch.relu(bn(conv(h, w0["conv2d_1/kernel"], 2, (2, 2, 2, 2)), w0["batch_normalization_1/beta"]))
    h = h.reshape(8, -1) @ w0["dense/kernel"]
    h = torch.relu(bn(h, w0["batch_normalization_2/beta"]))
    p = torch.softmax(h @ w0["dense_1/kernel"] + w0["dense_1/bias"], -1)
    loss = (-torch.log(p.gather(1, y).clamp(1e-7, 1 - 1e-7))).sum() / 8
    loss.backward()
    assert m.last_step_loss() == pytest.approx(loss.item(), rel=1e-4)
    for v in m.trainable_variables:
        np.testing.assert_allclose(v.grad.numpy(), w0[v.name].grad.numpy(), rtol=2e-3, atol=2e-6, err_msg=v.name)


def test_bn_moving_average_update_and_inference():
    x = torch.randn(64, 5) * 3 + 1
    mm, mv = torch.zeros(5), torch.ones(5)
    y, sm, si, nmm, nmv = ref.bn_fwd(x, None, torch.zeros(5), mm, mv, 0.99, 1e-3, True, False, False)
    np.testing.assert_allclose(nmm, 0.01 * x.mean(0), rtol=1e-5)
    np.testing.assert_allclose(nmv, 0.99 + 0.01 * x.var(0, unbiased=False), rtol=1e-5)
    _, _, _, _, nmv_u = ref.bn_fwd(x, None, torch.zeros(5), mm, mv, 0.99, 1e-3, True, False, True)
    np.testing.assert_allclose(nmv_u, 0.99 + 0.01 * x.var(0, unbiased=True), rtol=1e-5)
    yi = ref.bn_fwd(x, None, torch.zeros(5), nmm, nmv, training=False)[0]
    np.testing.assert_allclose(yi, (x - nmm) / torch.sqrt(nmv + 1e-3), rtol=1e-5, atol=1e-6)


def test_dropout_is_inverted_counter_based_and_identity_at_inference():
    x = torch.ones(10000)
    y = ref.dropout_fwd(x, 0.5, seed=42, step=3, stream=1)
    kept = (y != 0).float().mean().item()
    assert 0.47 < kept < 0.53 and set(y.unique().tolist()) == {0.0, 2.0}
    assert torch.equal(y, ref.dropout_fwd(x, 0.5, 42, 3, 1)), "same (seed, step, stream) -> same mask"
    assert not torch.equal(y, ref.dropout_fwd(x, 0.5, 42, 4, 1)), "mask changes with the step"
    assert not torch.equal(y, ref.dropout_fwd(x, 0.5, 42, 3, 2)), "mask changes with the stream"


def test_philox_known_answer():
    # Philox4x32-10 with zero counter/key (Random123 known-answer test vector)
    z = torch.zeros(1, dtype=torch.int64)
    out = ref.philox4x32_safe(z, z, z, z, 0, 0)
    assert [int(o) for o in out] == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]


def test_losses_logits_vs_probs_and_label_shapes():
    z = torch.randn(12, 10)
    y = torch.randint(0, 10, (12,))
    l1, c1, d1 = ref.xent_logits(z, y, 1 / 12)
    l2, c2, d2 = ref.xent_probs(torch.softmax(z, -1), y.reshape(-1, 1), 1 / 12)
    assert l1.item() == pytest.approx(F.cross_entropy(z, y).item(), rel=1e-6)
    assert l2.item() == pytest.approx(l1.item(), rel=1e-5) and c1 == c2
    obj = tf.keras.losses.SparseCategoricalCrossentropy(from_logits=True, reduction=tf.keras.losses.Reduction.NONE)
    per = obj(y, z)
    assert per.shape == (12,) and per.sum().item() / 12 == pytest.approx(l1.item(), rel=1e-6)


@pytest.mark.parametrize("opt", ["sgd", "momentum", "nesterov", "adam"])
def test_optimizers_match_torch(opt):
    from tfde_b200.parallel.comm import Arena
    O = tf.keras.optimizers
    n = 50
    w0 = torch.randn(n)
    mk = {"sgd": lambda: O.SGD(0.1), "momentum": lambda: O.SGD(0.1, momentum=0.9),
          "nesterov": lambda: O.SGD(0.1, momentum=0.9, nesterov=True), "adam": lambda: O.Adam(0.01, epsilon=1e-7)}[opt]
    o = mk().bind("cpu")
    a = Arena(n, "cpu")
    a.w[:n] = w0
    p = torch.nn.Parameter(w0.clone())
    t = {"sgd": lambda: torch.optim.SGD([p], 0.1), "momentum": lambda: torch.optim.SGD([p], 0.1, momentum=0.9),
         "nesterov": lambda: torch.optim.SGD([p], 0.1, momentum=0.9, nesterov=True),
         "adam": lambda: torch.optim.Adam([p], 0.01, eps=1e-7)}[opt]()
    for i in range(4):
        g = torch.randn(n, generator=torch.Generator().manual_seed(i))
        a.g[:n] = g
        o.apply_local(a)
        p.grad = g.clone()
        t.step()
    tol = 1e-5 if opt != "adam" else 2e-4   # Keras' epsilon placement differs from torch's by O(eps)
    np.testing.assert_allclose(a.w[:n].numpy(), p.detach().numpy(), rtol=tol, atol=tol)


def test_fit_history_evaluate_predict_and_weights_roundtrip(tmp_path):
    (xtr, ytr), (xte, yte) = tf.data.mnist.load_data(n_train=512, n_test=128)
    with cpu_scope():
        m = tf.models.build_and_compile_cnn_model(learning_rate=0.05, seed=1)
    ds = tf.data.Dataset.from_tensor_slices((xtr[..., None], ytr.astype(np.int64))).shuffle(512, seed=0).repeat().batch(32)
    h = m.fit(ds, epochs=2, steps_per_epoch=8, verbose=0)
    assert len(h.history["loss"]) == 2 and h.history["loss"][1] < h.history["loss"][0]
    loss, acc = m.evaluate(xte[..., None], yte.astype(np.int64), batch_size=50)
    assert 0 <= acc <= 1 and loss > 0
    pred = m.predict(xte[:10, ..., None])
    assert pred.shape == (10, 10)
    p = str(tmp_path / "w.npz")
    m.save_weights(p)
    with cpu_scope():
        m2 = tf.models.build_and_compile_cnn_model(seed=99)
        m2.build()
    m2.load_weights(p)
    np.testing.assert_array_equal(m2.predict(xte[:10, ..., None]), pred)
    ws = m.get_weights()
    m2.set_weights(ws)
    assert all(np.array_equal(a, b) for a, b in zip(m2.get_weights(), ws))
