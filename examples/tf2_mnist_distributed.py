"""The B200-native equivalent of the reference's tf2_mnist_distributed.py: module constants instead of
flags, ``ParameterServerStrategy`` handed to ``RunConfig(train_distribute=...)``, Model B as an Estimator,
``train_and_evaluate``; plus the hand-written ``model_fn`` around Model A (dead code in the reference,
kept runnable here behind ``--model-fn``)."""
import logging
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np

import tfde_b200 as tf
from tfde_b200 import estimator as tfe

BATCH_SIZE = 128
BUFFER_SIZE = 10000
LEARNING_RATE = 1e-4
MODEL_DIR = os.environ.get("TFDE_MODEL_DIR", "/tmp/mode")


def input_fn(features, labels, batch_size, mode):
    inputs = features if labels is None else (features, labels)
    dataset = tf.data.Dataset.from_tensor_slices(inputs)
    if mode == tfe.ModeKeys.TRAIN:
        dataset = dataset.shuffle(1000).repeat().batch(batch_size)
        dataset = dataset.prefetch(100)
    if mode in (tfe.ModeKeys.EVAL, tfe.ModeKeys.PREDICT):
        dataset = dataset.batch(batch_size)
    return dataset


def model_fn(features, labels, mode):
    model = tf.keras.Sequential([
        tf.keras.layers.Conv2D(32, 3, activation='relu', input_shape=(28, 28, 1)),
        tf.keras.layers.MaxPooling2D(),
        tf.keras.layers.Flatten(),
        tf.keras.layers.Dense(64, activation='relu'),
        tf.keras.layers.Dense(10)
    ])
    logits = model(features, training=False)
    if mode == tfe.ModeKeys.PREDICT:
        return tfe.EstimatorSpec(mode, predictions={'logits': logits})
    optimizer = tf.keras.optimizers.SGD(learning_rate=LEARNING_RATE)
    loss = tf.keras.losses.SparseCategoricalCrossentropy(
        from_logits=True, reduction=tf.keras.losses.Reduction.NONE)(labels, logits)
    loss = tfe.reduce_sum(loss) * (1. / BATCH_SIZE)
    if mode == tfe.ModeKeys.EVAL:
        return tfe.EstimatorSpec(mode, loss=loss)
    return tfe.EstimatorSpec(mode=mode, loss=loss,
                             train_op=optimizer.minimize(loss, tfe.get_or_create_global_step()))


def create_model(model_dir, config, learning_rate):
    def build():
        model = tf.models.build_model_b()
        optimizer = tf.keras.optimizers.SGD(learning_rate)   # the reference ignores its argument here
        model.compile(optimizer=optimizer, loss='sparse_categorical_crossentropy', metrics=['accuracy'])
        return model
    tf.keras.backend.set_learning_phase(True)
    if config.is_chief and config.task_type != 'ps':
        build().summary()
    return tfe.model_to_estimator(keras_model_builder=build, model_dir=model_dir, config=config)


def serving_input_fn():
    feature_placeholder = tfe.placeholder('float32', [None, 28 * 28])
    return tfe.export.TensorServingInputReceiver(feature_placeholder, feature_placeholder)


def main():
    logging.getLogger().setLevel(logging.INFO)
    tf.obs.set_verbosity("INFO")
    strategy = tf.distribute.experimental.ParameterServerStrategy()
    (train_images, train_labels), (test_images, test_labels) = tf.keras.datasets.mnist.load_data()
    train_images = train_images / 255.0
    test_images = test_images / 255.0
    train_labels = np.asarray(train_labels).astype('int').reshape((-1, 1))
    test_labels = np.asarray(test_labels).astype('int').reshape((-1, 1))
    train_steps = float(os.environ.get("TFDE_MAX_STEPS", len(train_images) / BATCH_SIZE))

    run_config = tfe.RunConfig(train_distribute=strategy)
    if "--model-fn" in sys.argv:
        estimator = tfe.Estimator(model_fn=model_fn, model_dir=MODEL_DIR, config=run_config)
        tr = lambda: input_fn(train_images.reshape(-1, 28, 28, 1), train_labels, BATCH_SIZE, tfe.ModeKeys.TRAIN)
        ev = lambda: input_fn(test_images.reshape(-1, 28, 28, 1), test_labels, BATCH_SIZE, tfe.ModeKeys.EVAL)
    else:
        estimator = create_model(model_dir=MODEL_DIR, config=run_config, learning_rate=LEARNING_RATE)
        tr = lambda: input_fn(train_images, train_labels, BATCH_SIZE, tfe.ModeKeys.TRAIN)
        ev = lambda: input_fn(test_images, test_labels, BATCH_SIZE, tfe.ModeKeys.EVAL)
    train_spec = tfe.TrainSpec(input_fn=tr, max_steps=train_steps)
    exporter = tfe.FinalExporter('exporter', serving_input_fn)
    eval_spec = tfe.EvalSpec(input_fn=ev, steps=None, name='mnist-eval', exporters=[exporter],
                             start_delay_secs=10, throttle_secs=10)
    return tfe.train_and_evaluate(estimator, train_spec, eval_spec)


if __name__ == '__main__':
    main()
