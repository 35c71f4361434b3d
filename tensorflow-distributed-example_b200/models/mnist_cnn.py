"""The two MNIST classifiers of the reference, built with this framework's Keras-like API.

* Model A -- ``build_and_compile_cnn_model`` (distributed_with_keras.py:32-44; duplicate in the unused
  ``model_fn`` of tf2_mnist_distributed.py:66-72): Conv32(k3,relu) - MaxPool - Flatten - Dense64(relu) -
  Dense10 logits; SCCE(from_logits) + SGD(0.001) + accuracy. 347,146 parameters.
* Model B -- ``create_model`` (mnist_keras_distributed.py:79-109, tf2_mnist_distributed.py:105-135):
  Reshape - 3x(Conv-BN-ReLU) - Flatten - Dense200-BN-ReLU - Dropout(0.5) - Dense10 softmax;
  'sparse_categorical_crossentropy' on probabilities + SGD + accuracy. 250,466 trainable + 484
  non-trainable parameters.
"""
from __future__ import annotations

from ..keras import layers as l
from ..keras import losses, optimizers
from ..keras.models import Sequential

MODEL_A_PARAMS = 347_146
MODEL_B_TRAINABLE = 250_466
MODEL_B_NON_TRAINABLE = 484


def build_model_a(seed=None) -> Sequential:
    return Sequential([
        l.Conv2D(32, 3, activation="relu", input_shape=(28, 28, 1)),
        l.MaxPooling2D(),
        l.Flatten(),
        l.Dense(64, activation="relu"),
        l.Dense(10),
    ], seed=seed)


def build_and_compile_cnn_model(learning_rate=0.001, optimizer=None, seed=None) -> Sequential:
    model = build_model_a(seed=seed)
    model.compile(loss=losses.SparseCategoricalCrossentropy(from_logits=True),
                  optimizer=optimizer or optimizers.SGD(learning_rate=learning_rate),
                  metrics=["accuracy"])
    return model


def build_model_b(seed=None) -> Sequential:
    return Sequential([
        l.Reshape(input_shape=(28 * 28,), target_shape=(28, 28, 1)),
        l.Conv2D(filters=6, kernel_size=3, padding="same", use_bias=False),
        l.BatchNormalization(scale=False, center=True),
        l.Activation("relu"),
        l.Conv2D(filters=12, kernel_size=6, padding="same", use_bias=False, strides=2),
        l.BatchNormalization(scale=False, center=True),
        l.Activation("relu"),
        l.Conv2D(filters=24, kernel_size=6, padding="same", use_bias=False, strides=2),
        l.BatchNormalization(scale=False, center=True),
        l.Activation("relu"),
        l.Flatten(),
        l.Dense(200, use_bias=False),
        l.BatchNormalization(scale=False, center=True),
        l.Activation("relu"),
        l.Dropout(0.5),
        l.Dense(10, activation="softmax"),
    ], seed=seed)


def build_and_compile_model_b(learning_rate=0.01, optimizer=None, seed=None) -> Sequential:
    model = build_model_b(seed=seed)
    model.compile(optimizer=optimizer or optimizers.GradientDescentOptimizer(learning_rate),
                  loss="sparse_categorical_crossentropy", metrics=["accuracy"])
    return model
