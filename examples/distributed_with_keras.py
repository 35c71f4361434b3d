"""Synchronous multi-worker training of the Keras MNIST CNN -- the B200-native equivalent of the
reference's distributed_with_keras.py (same constants, same call sequence, same TF_CONFIG contract).

    python -m tfde_b200.launch --nproc 2 examples/distributed_with_keras.py      # 2 workers, this host
    TF_CONFIG='{"cluster": {"worker": ["h1:port", "h2:port"]}, "task": {"type": "worker", "index": 0}}' \
        python examples/distributed_with_keras.py                                 # one shell per worker
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np

import tfde_b200 as tf
from tfde_b200.data import mnist as tfds

tfds.disable_progress_bar()

BUFFER_SIZE = 10000
BATCH_SIZE = 64
strategy = tf.distribute.experimental.MultiWorkerMirroredStrategy()
NUM_WORKERS = strategy.num_replicas_in_sync          # the reference hard-codes 2
GLOBAL_BATCH_SIZE = BATCH_SIZE * NUM_WORKERS


def make_datasets_unbatched():
    # Scaling MNIST data from (0, 255] to (0., 1.]
    def scale(image, label):
        image = image.astype(np.float32)
        image /= 255
        return image, label

    datasets, info = tfds.load(name='mnist', data_dir='/tmp/data', with_info=True, as_supervised=True)
    return datasets['train'].map(scale).cache().shuffle(BUFFER_SIZE)


def build_and_compile_cnn_model():
    model = tf.keras.Sequential([
        tf.keras.layers.Conv2D(32, 3, activation='relu', input_shape=(28, 28, 1)),
        tf.keras.layers.MaxPooling2D(),
        tf.keras.layers.Flatten(),
        tf.keras.layers.Dense(64, activation='relu'),
        tf.keras.layers.Dense(10)
    ])
    model.compile(
        loss=tf.keras.losses.SparseCategoricalCrossentropy(from_logits=True),
        optimizer=tf.keras.optimizers.SGD(learning_rate=0.001),
        metrics=['accuracy'])
    return model


def main():
    with strategy.scope():
        # Creation of dataset, and model building/compiling need to be within `strategy.scope()`.
        train_datasets = make_datasets_unbatched().batch(GLOBAL_BATCH_SIZE)
        options = tf.data.Options()
        options.experimental_distribute.auto_shard_policy = tf.data.experimental.AutoShardPolicy.OFF
        train_datasets_no_auto_shard = train_datasets.with_options(options)
        multi_worker_model = build_and_compile_cnn_model()

    history = multi_worker_model.fit(x=train_datasets_no_auto_shard, epochs=3, steps_per_epoch=5)
    strategy.shutdown()
    return history


if __name__ == '__main__':
    main()
