"""tfde_b200 -- a B200-native synchronous data-parallel training framework with the capabilities of
lowc1012/tensorflow-distributed-example (tf.distribute strategies, tf.data-like input, Keras-like
models, Estimator-like train_and_evaluate), re-designed for Blackwell: hand-written sm_100a kernels,
CUDA-graph captured steps, and a fused peer-to-peer all-reduce+optimizer over NVLink/NVSwitch.

Namespace mirrors what the reference scripts import from TensorFlow::

    import tfde_b200 as tf
    strategy = tf.distribute.experimental.MultiWorkerMirroredStrategy()
    with strategy.scope():
        model = tf.keras.Sequential([...]); model.compile(...)
    model.fit(dataset, epochs=3, steps_per_epoch=5)
"""
__version__ = "0.1.0"

from . import ops  # noqa: F401
from . import keras  # noqa: F401
from . import data  # noqa: F401
from . import distribute  # noqa: F401
from . import cluster  # noqa: F401
from . import estimator  # noqa: F401
from . import ckpt  # noqa: F401
from . import obs  # noqa: F401
from . import models  # noqa: F401
from . import parallel  # noqa: F401
from . import utils  # noqa: F401
from .keras.models import set_seed  # noqa: F401


class random:  # tf.random.set_seed
    set_seed = staticmethod(set_seed)
