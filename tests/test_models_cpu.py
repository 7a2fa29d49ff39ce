"""The larger model families of BASELINE.json (ResNet-50, BERT-base MLM): topology, parameter counts, one
training step on CPU with tiny inputs."""
import numpy as np
import pytest

import tfdl
from tfdl import keras
from tfdl.models import BertMLM, ResNet50, build_and_compile_bert_mlm


def test_resnet50_parameter_count_matches_keras_applications():
    keras.clear_session()
    m = ResNet50(classes=1000, input_shape=(224, 224, 3))
    assert m.count_params() == 25_636_712
    assert sum(int(np.prod(w.shape)) for w in m.weights if not w.trainable) == 53_120
    names = [w.name for w in m.weights]
    assert names[0] == "conv1_conv/kernel:0" and "conv5_block3_3_bn/moving_variance:0" in names
    assert m.weights[0].shape == (7, 7, 3, 64) and names[-2:] == ["predictions/kernel:0", "predictions/bias:0"]


def test_resnet50_trains_a_step_on_small_images():
    keras.clear_session()
    m = ResNet50(classes=10, input_shape=(32, 32, 3))
    m.compile(optimizer=keras.optimizers.SGD(0.01, momentum=0.9),
              loss=keras.losses.SparseCategoricalCrossentropy(from_logits=True), metrics=["accuracy"])
    rng = np.random.default_rng(0)
    x, y = rng.random((4, 32, 32, 3), dtype=np.float32), rng.integers(0, 10, size=4)
    w0 = m.get_weights()[0].copy()
    l0 = m.train_on_batch(x, y, return_dict=True)["loss"]
    l1 = m.train_on_batch(x, y, return_dict=True)["loss"]
    assert np.isfinite(l0) and np.isfinite(l1) and not np.array_equal(w0, m.get_weights()[0])
    mm = [w for w in m.weights if w.name == "conv1_bn/moving_mean:0"][0]
    assert np.abs(mm.numpy()).max() > 0  # BatchNorm statistics are tracked but not touched by the optimizer


def test_bert_mlm_small_config_trains_and_counts():
    keras.clear_session()
    m = BertMLM(vocab_size=1000, hidden=64, layers=2, heads=4, intermediate=128, max_position=32, seq_len=16)
    m.compile(optimizer=keras.optimizers.Adam(1e-3),
              loss=keras.losses.SparseCategoricalCrossentropy(from_logits=True, ignore_class=-100))
    per_layer = 4 * (64 * 64 + 64) + 2 * 64 * 2 + (64 * 128 + 128) + (128 * 64 + 64)
    expect = 1000 * 64 + 32 * 64 + 2 * 64 + 2 * 64 + 2 * per_layer + (64 * 64 + 64) + 2 * 64 + 1000
    assert m.count_params() == expect
    rng = np.random.default_rng(0)
    ids = rng.integers(0, 1000, size=(4, 16))
    labels = np.where(rng.random((4, 16)) < 0.15, ids, -100)
    labels[:, 0] = ids[:, 0]
    losses = [m.train_on_batch(ids, labels, return_dict=True)["loss"] for _ in range(5)]
    assert all(np.isfinite(losses)) and losses[-1] < losses[0]


def test_bert_base_parameter_count():
    # count from the shapes only (building 110 M random parameters is too slow for a unit test)
    H, V, P, I, Lr = 768, 30522, 512, 3072, 12
    per_layer = 4 * (H * H + H) + 2 * 2 * H + (H * I + I) + (I * H + H)
    total = V * H + P * H + 2 * H + 2 * H + Lr * per_layer + (H * H + H) + 2 * H + V
    assert 109_000_000 < total < 111_000_000
