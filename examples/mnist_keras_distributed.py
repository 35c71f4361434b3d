"""Runs MNIST using Distribution Strategies -- the B200-native equivalent of the reference's
mnist_keras_distributed.py: same flags, same env contract (CLUSTER_SPEC / JOB_NAME / TASK_INDEX -> TF_CONFIG,
TB_PORT), same flow (RunConfig with ParameterServerStrategy for training and MirroredStrategy for
evaluation, Keras Model B -> Estimator, TrainSpec/EvalSpec/FinalExporter, TensorBoard on worker 0,
train_and_evaluate).

    python examples/mnist_keras_distributed.py --working-dir /tmp/mnist                   # local
    python -m tfde_b200.launch --nproc 2 --ps 1 --env-style cluster_spec \
        examples/mnist_keras_distributed.py --working-dir /tmp/mnist                      # async PS job

Differences from the reference, on purpose: --num-epochs is honoured
(max_steps = epochs * len(train) / batch), job_type/job_index are always bound, `== 0` instead of `is 0`.
"""
import argparse
import logging
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np

import tfde_b200 as tf
from tfde_b200 import estimator as tfe
from tfde_b200.obs import set_verbosity, start_tensorboard


def get_args():
    parser = argparse.ArgumentParser()
    parser.add_argument('--working-dir', type=str, required=True,
                        help='location to write checkpoints and export models')
    parser.add_argument('--num-epochs', type=float, default=5,
                        help='number of times to go through the data, default=5')
    parser.add_argument('--batch-size', default=128, type=int,
                        help='number of records to read during each training step, default=128')
    parser.add_argument('--learning-rate', default=.01, type=float,
                        help='learning rate for gradient descent, default=.01')
    parser.add_argument('--verbosity', choices=['DEBUG', 'ERROR', 'FATAL', 'INFO', 'WARN'], default='INFO')
    parser.add_argument('--max-steps', type=float, default=None,
                        help='override the number of training steps (default: num_epochs * len(train) / batch)')
    args, _ = parser.parse_known_args()
    return args


def create_model(model_dir, config, learning_rate):
    """Creates the Keras Sequential model (3x Conv-BN-ReLU, Dense200-BN-ReLU-Dropout, Dense10 softmax) and
    turns it into an Estimator."""
    l = tf.keras.layers

    def build():
        model = tf.keras.Sequential([
            l.Reshape(input_shape=(28 * 28,), target_shape=(28, 28, 1)),
            l.Conv2D(filters=6, kernel_size=3, padding='same', use_bias=False),  # no bias necessary before batch norm
            l.BatchNormalization(scale=False, center=True),  # no batch norm scaling necessary before "relu"
            l.Activation('relu'),  # activation after batch norm
            l.Conv2D(filters=12, kernel_size=6, padding='same', use_bias=False, strides=2),
            l.BatchNormalization(scale=False, center=True),
            l.Activation('relu'),
            l.Conv2D(filters=24, kernel_size=6, padding='same', use_bias=False, strides=2),
            l.BatchNormalization(scale=False, center=True),
            l.Activation('relu'),
            l.Flatten(),
            l.Dense(200, use_bias=False),
            l.BatchNormalization(scale=False, center=True),
            l.Activation('relu'),
            l.Dropout(0.5),  # Dropout on dense layer only
            l.Dense(10, activation='softmax')
        ])
        optimizer = tf.keras.optimizers.GradientDescentOptimizer(learning_rate)
        model.compile(optimizer=optimizer, loss='sparse_categorical_crossentropy', metrics=['accuracy'])
        return model

    tf.keras.backend.set_learning_phase(True)
    if config.is_chief and config.task_type != 'ps':
        build().summary()
    return tfe.model_to_estimator(keras_model_builder=build, model_dir=model_dir, config=config)


def input_fn(features, labels, batch_size, mode):
    if labels is None:
        inputs = features
    else:
        inputs = (features, labels)
    dataset = tf.data.Dataset.from_tensor_slices(inputs)
    if mode == tfe.ModeKeys.TRAIN:
        dataset = dataset.shuffle(1000).repeat().batch(batch_size)
        dataset = dataset.prefetch(100)
    if mode in (tfe.ModeKeys.EVAL, tfe.ModeKeys.PREDICT):
        dataset = dataset.batch(batch_size)
    return dataset


def serving_input_fn():
    """Defines the features to be passed to the model during inference: flat 784-float rows."""
    feature_placeholder = tfe.placeholder('float32', [None, 28 * 28])
    features = feature_placeholder
    return tfe.export.TensorServingInputReceiver(features, feature_placeholder)


def train_and_evaluate(args):
    (train_images, train_labels), (test_images, test_labels) = tf.keras.datasets.mnist.load_data()
    train_images = train_images / 255.0
    test_images = test_images / 255.0
    train_labels = np.asarray(train_labels).astype('int').reshape((-1, 1))
    test_labels = np.asarray(test_labels).astype('int').reshape((-1, 1))

    train_steps = args.max_steps if args.max_steps is not None else \
        args.num_epochs * len(train_images) / args.batch_size

    job_type, job_index = 'worker', 0
    if tf.cluster.synthesize_tf_config_from_cluster_spec() is not None:
        job_index = int(os.environ["TASK_INDEX"])
        job_type = os.environ["JOB_NAME"]

    # hook = tfe.ProfilerHook(save_steps=100, output_dir=args.working_dir, show_memory=True)

    run_config = tfe.RunConfig(
        experimental_distribute=tf.distribute.DistributeConfig(
            train_distribute=tf.distribute.experimental.ParameterServerStrategy(),
            eval_distribute=tf.distribute.MirroredStrategy()),
        session_config=tfe.session_config_from_env(),
        model_dir=args.working_dir,
        save_summary_steps=100,
        log_step_count_steps=100,
        save_checkpoints_steps=500)
    estimator = create_model(model_dir=args.working_dir, config=run_config, learning_rate=args.learning_rate)
    train_spec = tfe.TrainSpec(
        input_fn=lambda: input_fn(train_images, train_labels, args.batch_size, mode=tfe.ModeKeys.TRAIN),
        # hooks=[hook],  # Uncomment if needed to debug.
        max_steps=train_steps)
    exporter = tfe.FinalExporter('exporter', serving_input_fn)
    eval_spec = tfe.EvalSpec(
        input_fn=lambda: input_fn(test_images, test_labels, args.batch_size, mode=tfe.ModeKeys.EVAL),
        steps=None, name='mnist-eval', exporters=[exporter], start_delay_secs=10, throttle_secs=10)

    if job_type == "worker" and job_index == 0 and os.environ.get("TFDE_NO_TENSORBOARD") != "1":
        os.makedirs(args.working_dir, exist_ok=True)
        start_tensorboard(args.working_dir)

    return tfe.train_and_evaluate(estimator, train_spec, eval_spec)


if __name__ == '__main__':
    args = get_args()
    set_verbosity(args.verbosity)
    logging.getLogger().setLevel(logging.INFO)
    train_and_evaluate(args)
