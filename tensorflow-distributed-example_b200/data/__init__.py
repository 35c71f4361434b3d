"""tf.data surface."""
from .dataset import Dataset, Options, AutoShardPolicy, DistributedDataset  # noqa: F401
from . import mnist  # noqa: F401


class experimental:
    AutoShardPolicy = AutoShardPolicy
