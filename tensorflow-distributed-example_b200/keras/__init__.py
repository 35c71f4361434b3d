"""tf.keras surface: Sequential, layers, losses, optimizers, callbacks, datasets."""
from . import layers, losses, optimizers, callbacks  # noqa: F401
from .models import Sequential, set_seed  # noqa: F401


class datasets:  # tf.keras.datasets.mnist.load_data()
    class mnist:
        @staticmethod
        def load_data(*a, **k):
            from ..data.mnist import load_data
            return load_data(*a, **k)


class backend:
    _learning_phase = None

    @staticmethod
    def set_learning_phase(value):
        """tf.keras.backend.set_learning_phase (mnist_keras_distributed.py:116). Training/inference is an
        explicit argument of every forward here, so this only records the value."""
        backend._learning_phase = value


class mixed_precision:
    """tf.keras.mixed_precision: 'mixed_bfloat16' routes large Dense layers to the tcgen05 bf16 GEMM
    (fp32 master weights, fp32 accumulation); 'float32' (default) is the reference's precision."""

    @staticmethod
    def set_global_policy(name, activations=None):
        """``activations='bfloat16'`` (extension): bf16 layer outputs as in Keras' own mixed precision
        (convolutional networks); default keeps fp32 activations and bf16 GEMM operands only."""
        from ..ops import mixed
        mixed.set_global_policy(name, activations)

    @staticmethod
    def activation_dtype():
        from ..ops import mixed
        return mixed.activation_dtype()

    @staticmethod
    def global_policy():
        from ..ops import mixed
        return mixed.global_policy()
