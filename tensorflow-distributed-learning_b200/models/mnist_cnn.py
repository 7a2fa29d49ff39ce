"""The reference model, line for line in this framework's API
(/root/reference/tf_dist_example.py:39-53): Conv32-Pool-Conv64-Pool-Flatten-Dense128-Dense10,
SparseCategoricalCrossentropy(from_logits=True), SGD(lr=0.001), SparseCategoricalAccuracy.
225,034 parameters in 8 variables."""
from __future__ import annotations

from .. import keras


def build_cnn_model() -> "keras.Sequential":
    L = keras.layers
    return keras.Sequential([
        L.Conv2D(32, 3, activation="relu", input_shape=(28, 28, 1)),
        L.MaxPooling2D(),
        L.Conv2D(64, 3, activation="relu"),
        L.MaxPooling2D(),
        L.Flatten(),
        L.Dense(128, activation="relu"),
        L.Dense(10),
    ])


def build_and_compile_cnn_model(learning_rate: float = 0.001, momentum: float = 0.0) -> "keras.Sequential":
    model = build_cnn_model()
    model.compile(loss=keras.losses.SparseCategoricalCrossentropy(from_logits=True),
                  optimizer=keras.optimizers.SGD(learning_rate=learning_rate, momentum=momentum),
                  metrics=[keras.metrics.SparseCategoricalAccuracy()])
    return model
