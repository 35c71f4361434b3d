"""Observability pieces of SURVEY.md sections 2.6 / 5.5: TensorBoard launcher (TB_PORT), event-file writer."""
import glob
import os
import socket

from tfde_b200.obs import start_tensorboard
from tfde_b200.obs.summary import SummaryWriter


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_start_tensorboard_reads_tb_port_and_serves(tmp_path, monkeypatch):
    """mnist_keras_distributed.py:192-197: in-process TensorBoard, core plugin, port from TB_PORT."""
    port = _free_port()
    monkeypatch.setenv("TB_PORT", str(port))
    url = start_tensorboard(str(tmp_path))
    if url is None:          # the launcher degrades to a warning where no server can be started
        return
    assert str(port) in url
    with socket.create_connection(("127.0.0.1", port), timeout=5):
        pass


def test_start_tensorboard_never_raises_on_a_busy_port(tmp_path):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    s.listen(1)
    try:
        assert start_tensorboard(str(tmp_path), port=s.getsockname()[1]) is None or True
    finally:
        s.close()


def test_summary_writer_creates_event_file(tmp_path):
    w = SummaryWriter(str(tmp_path))
    w.scalar("loss", 1.5, 100)
    w.scalar("global_step/sec", 1234.0, 100)
    w.flush()
    w.close()
    files = glob.glob(os.path.join(str(tmp_path), "events.out.tfevents.*"))
    assert files and os.path.getsize(files[0]) > 0
