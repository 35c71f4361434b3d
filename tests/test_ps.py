"""Parameter-server transport (parallel/ps.py) on CPU: the pickle-free wire format, the listener's key policy, and
Adam on several ps tasks (the bias correction must not depend on which ps task owns a variable)."""
import json
import socket
import struct
import threading

import numpy as np
import pytest
import torch

import tfde_b200 as tf
from tfde_b200.cluster import tf_config as tfc
from tfde_b200.parallel import ps as P


def _free_ports(n):
    socks = [socket.socket() for _ in range(n)]
    for s in socks:
        s.bind(("127.0.0.1", 0))
    ports = [s.getsockname()[1] for s in socks]
    for s in socks:
        s.close()
    return ports


class _Pipe:
    def __init__(self):
        self.buf = None

    def send_bytes(self, b):
        self.buf = bytes(b)

    def recv_bytes(self, maxlength=None):
        return self.buf


def test_wire_round_trip_and_rejections():
    c = _Pipe()
    a = {"w": np.arange(6, dtype=np.float32).reshape(2, 3), "y": np.array([1, 2], dtype=np.int64)}
    P._send(c, "push", {"lr": 0.5}, a)
    op, meta, arr = P._recv(c)
    assert op == "push" and meta == {"lr": 0.5}
    np.testing.assert_array_equal(arr["w"], a["w"])
    np.testing.assert_array_equal(arr["y"], a["y"])
    with pytest.raises(TypeError):
        P._send(c, "push", {}, {"o": np.array([object()], dtype=object)})     # nothing that needs pickle
    # a forged header naming a dtype outside the allow-list, and a payload shorter than its header claims
    hdr = json.dumps({"op": "x", "meta": {}, "arrays": [["a", "object", [1]]]}).encode()
    c.buf = struct.pack("<I", len(hdr)) + hdr + b"\x00" * 8
    with pytest.raises(ValueError):
        P._recv(c)
    hdr = json.dumps({"op": "x", "meta": {}, "arrays": [["a", "float32", [1000]]]}).encode()
    c.buf = struct.pack("<I", len(hdr)) + hdr + b"\x00" * 8
    with pytest.raises(ValueError):
        P._recv(c)
    # a pickle is just bytes here: it fails the header check instead of being executed
    import pickle
    c.buf = pickle.dumps(("push", {"w": a["w"]}))
    with pytest.raises(Exception):
        P._recv(c)


def test_default_key_is_refused_off_loopback(monkeypatch):
    monkeypatch.delenv("TFDE_PS_AUTHKEY", raising=False)
    monkeypatch.delenv("TFDE_PS_AUTHKEY_FILE", raising=False)
    with pytest.raises(RuntimeError, match="TFDE_PS_AUTHKEY"):
        P.ParameterServer("10.1.2.3", 1, 0).serve_forever()
    monkeypatch.setenv("TFDE_PS_AUTHKEY", "job-secret")
    assert P._authkey() == b"job-secret"


def _train_with_ps(n_ps, kind, steps=5):
    ports = _free_ports(n_ps + 1)
    cluster = {"worker": [f"127.0.0.1:{ports[0]}"], "ps": [f"127.0.0.1:{p}" for p in ports[1:]]}
    servers = [P.ParameterServer("127.0.0.1", p, i) for i, p in enumerate(ports[1:])]
    threads = [threading.Thread(target=s.serve_forever, daemon=True) for s in servers]
    for t in threads:
        t.start()
    res = tfc.parse_tf_config(json.dumps({"cluster": cluster, "task": {"type": "worker", "index": 0}}))
    st = tf.distribute.experimental.ParameterServerStrategy(cluster_resolver=res, device="cpu")
    opt = tf.keras.optimizers.Adam(0.01) if kind == "adam" else tf.keras.optimizers.SGD(0.05)
    with st.scope():
        m = tf.models.build_and_compile_cnn_model(optimizer=opt, seed=3)
    g = torch.Generator().manual_seed(0)
    x = torch.rand(steps, 16, 28, 28, 1, generator=g)
    y = torch.randint(0, 10, (steps, 16), generator=g)
    st.on_train_begin(m)
    for k in range(steps):
        m.train_on_batch(x[k], y[k])
    step = st.global_step()
    st.on_train_end(m)
    for t in threads:
        t.join(timeout=10)
    return m.get_weights(), step


@pytest.mark.parametrize("kind", ["sgd", "adam"])
def test_two_ps_tasks_equal_one_ps_task(kind):
    """Round-robin placement over 2 ps tasks must train exactly like 1 ps task (Adam on ps 1..n-1 used
    ps 0's step for its bias correction and stayed at t = 1)."""
    w1, s1 = _train_with_ps(1, kind)
    w2, s2 = _train_with_ps(2, kind)
    assert s1 == s2 == 5
    for a, b in zip(w1, w2):
        np.testing.assert_allclose(a, b, rtol=1e-6, atol=1e-7)
    # and Adam on the ps equals the local Adam of the same model (single worker => no asynchrony)
    opt = tf.keras.optimizers.Adam(0.01) if kind == "adam" else tf.keras.optimizers.SGD(0.05)
    from tfde_b200.parallel.strategy import Strategy
    with Strategy(device="cpu").scope():
        m = tf.models.build_and_compile_cnn_model(optimizer=opt, seed=3)
    g = torch.Generator().manual_seed(0)
    x = torch.rand(5, 16, 28, 28, 1, generator=g)
    y = torch.randint(0, 10, (5, 16), generator=g)
    for k in range(5):
        m.train_on_batch(x[k], y[k])
    for a, b in zip(w2, m.get_weights()):
        np.testing.assert_allclose(a, b, rtol=2e-4, atol=2e-6)
