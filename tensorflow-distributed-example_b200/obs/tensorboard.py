"""In-process TensorBoard launcher (mnist_keras_distributed.py:192-197): core plugin only, port from
``TB_PORT`` (default 6006), background thread, INFO log line. Chief/worker-0 only at the call site
(mnist_keras_distributed.py:277-280)."""
from __future__ import annotations

import logging
import os


def start_tensorboard(logdir: str, port: int | None = None, host: str = "127.0.0.1"):
    """Returns the URL, or None when the tensorboard package cannot be launched here."""
    port = int(os.getenv("TB_PORT", 6006)) if port is None else int(port)
    try:
        from tensorboard import program as tb_program
        from tensorboard.plugins.core import core_plugin
        tb = tb_program.TensorBoard(plugins=[core_plugin.CorePluginLoader()])
        tb.configure(logdir=logdir, port=port, host=host)
        url = tb.launch()
    except Exception as e:  # no network stack / port in use: training must not die because of the UI
        logging.warning("TensorBoard not started (%s)", e)
        return None
    logging.info("Starting TensorBoard with --logdir=%s" % logdir)
    return url
