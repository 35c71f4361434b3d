"""tf.distribute surface."""
from .parallel.strategy import (Strategy, MirroredStrategy, MultiWorkerMirroredStrategy,  # noqa: F401
                                OneDeviceStrategy, get_strategy, has_strategy)
from .parallel.ps import ParameterServerStrategy  # noqa: F401


class experimental:
    MultiWorkerMirroredStrategy = MultiWorkerMirroredStrategy
    ParameterServerStrategy = ParameterServerStrategy


class DistributeConfig:
    """tf.contrib.distribute.DistributeConfig(train_distribute, eval_distribute)
    (mnist_keras_distributed.py:241-243)."""

    def __init__(self, train_distribute=None, eval_distribute=None, remote_cluster=None):
        self.train_distribute = train_distribute
        self.eval_distribute = eval_distribute
        self.remote_cluster = remote_cluster
