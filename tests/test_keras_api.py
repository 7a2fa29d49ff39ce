"""Keras call forms beyond the reference script's (`tf_dist_example.py`) that a user switching over will try first:
string optimizers / losses, RMSprop, binary cross-entropy, ReduceLROnPlateau, test_on_batch, keras.Input, utils."""
import numpy as np
import pytest
import torch

import tfdl
from tfdl import keras


def _xy(n=96, seed=0):
    rng = np.random.default_rng(seed)
    return rng.random((n, 28, 28, 1), dtype=np.float32), rng.integers(0, 10, size=n)


def _small(last=10):
    return keras.Sequential([keras.Input(shape=(28, 28, 1)), keras.layers.Conv2D(4, 3, activation="relu"),
                             keras.layers.MaxPooling2D(), keras.layers.Flatten(), keras.layers.Dense(last)])


@pytest.mark.parametrize("momentum", [0.0, 0.9])
def test_rmsprop_matches_torch(momentum):
    rng = np.random.default_rng(1)
    w0 = rng.standard_normal(50).astype(np.float32)
    grads = [rng.standard_normal(50).astype(np.float32) for _ in range(5)]
    opt = keras.optimizers.RMSprop(learning_rate=0.01, rho=0.8, momentum=momentum, epsilon=1e-6)
    params = torch.tensor(w0.copy())
    opt.bind([params])
    ref_w = torch.tensor(w0.copy(), requires_grad=True)
    ref = torch.optim.RMSprop([ref_w], lr=0.01, alpha=0.8, eps=1e-6, momentum=momentum)
    for g in grads:
        opt.apply_flat(params, torch.tensor(g), 0, None)
        opt.step_done()
        ref_w.grad = torch.tensor(g)
        ref.step()
    torch.testing.assert_close(params, ref_w.detach(), rtol=1e-5, atol=1e-6)
    assert keras.optimizers.from_config(opt.get_config()).get_config() == opt.get_config()


def test_string_forms_rmsprop_training_and_batch_methods():
    x, y = _xy()
    m = _small()
    assert len(m.layers) == 4  # the Input layer is not a layer of the model, as in Keras
    m.compile(optimizer="rmsprop", loss=keras.losses.SparseCategoricalCrossentropy(from_logits=True), metrics=["accuracy"])
    h = m.fit(x, y, verbose=0, epochs=4, workers=2)  # Keras-only tuning arguments are accepted
    assert h.history["loss"][-1] < h.history["loss"][0]
    assert [t.name for t in m.metrics] == ["loss", "accuracy"]
    before = [w.copy() for w in m.get_weights()]
    loss, acc = m.test_on_batch(x[:32], y[:32])
    want = m.evaluate(x[:32], y[:32], batch_size=32, return_dict=True)
    assert abs(loss - want["loss"]) < 1e-6 and abs(acc - want["accuracy"]) < 1e-6
    for a, b in zip(before, m.get_weights()):
        np.testing.assert_array_equal(a, b)  # no training happened
    assert m.predict_on_batch(x[:7]).shape == (7, 10)
    with pytest.raises(NotImplementedError, match="class_weight"):
        m.fit(x, y, verbose=0, class_weight={0: 2.0})
    with pytest.raises(ValueError, match="unknown optimizer"):
        m.compile(optimizer="lion", loss="mse")


def test_binary_crossentropy_and_accuracy():
    x, y = _xy()
    yb = (y > 4).astype(np.float32)
    m = keras.Sequential([keras.layers.Flatten(input_shape=(28, 28, 1)), keras.layers.Dense(1)])
    m.compile(optimizer=keras.optimizers.Adam(0.01), loss=keras.losses.BinaryCrossentropy(from_logits=True), metrics=["accuracy"])
    m.fit(x, yb, verbose=0, epochs=1)
    logits = torch.tensor(m.predict(x))
    got = m.evaluate(x, yb, batch_size=96, return_dict=True)
    want = torch.nn.functional.binary_cross_entropy_with_logits(logits, torch.tensor(yb).reshape(-1, 1)).item()
    assert abs(got["loss"] - want) < 1e-5
    assert abs(got["accuracy"] - float(((logits.reshape(-1) > 0).numpy() == (yb > 0.5)).mean())) < 1e-6
    p = keras.losses.BinaryCrossentropy()
    probs = torch.sigmoid(logits)
    assert abs(float(p(torch.tensor(yb), probs)) - want) < 1e-5
    assert abs(float(keras.losses.MeanAbsoluteError()(torch.tensor(yb), probs)) - float((probs.reshape(-1) - torch.tensor(yb)).abs().mean())) < 1e-6


def test_reduce_lr_on_plateau_and_full_model_checkpoints(tmp_path):
    x, y = _xy()
    m = _small()
    m.compile(optimizer=keras.optimizers.SGD(0.1), loss=keras.losses.SparseCategoricalCrossentropy(from_logits=True))
    cb = keras.callbacks.ReduceLROnPlateau(monitor="loss", factor=0.5, patience=1, min_delta=100.0, min_lr=0.02)
    ck = keras.callbacks.ModelCheckpoint(str(tmp_path / "full-{epoch}"), save_weights_only=False)
    h = m.fit(x, y, verbose=0, epochs=5, callbacks=[cb, ck])
    # nothing ever improves by 100: halved after every epoch but the first, down to the floor
    assert h.history["lr"] == pytest.approx([0.1, 0.1, 0.05, 0.025, 0.02])
    assert m.optimizer.learning_rate == pytest.approx(0.02)
    twin = keras.models.load_model(str(tmp_path / "full-5"))
    assert twin.optimizer.learning_rate == pytest.approx(0.02)
    np.testing.assert_allclose(twin.predict(x[:4]), m.predict(x[:4]), rtol=1e-6)


def test_utils():
    onehot = keras.utils.to_categorical([0, 2, 1], 4)
    assert onehot.shape == (3, 4) and onehot.dtype == np.float32 and onehot.argmax(1).tolist() == [0, 2, 1]
    keras.utils.set_random_seed(11)
    a = keras.layers.glorot_uniform((3, 3))
    keras.utils.set_random_seed(11)
    b = keras.layers.glorot_uniform((3, 3))
    tfdl.config.set_random_seed(None)
    np.testing.assert_array_equal(a, b)


def test_data_and_distribute_call_forms():
    """tf.data / tf.distribute spellings around the ones the reference uses: AUTOTUNE, filter / zip / reduce,
    MirroredStrategy(cross_device_ops=NcclAllReduce()) (the default the reference's README names, README.md:17),
    MultiWorkerMirroredStrategy(communication_options=...)."""
    D, T = tfdl.data, tfdl.distribute
    x = np.arange(40, dtype=np.float32).reshape(20, 2)
    y = np.arange(20) % 3
    ds = D.Dataset.from_tensor_slices((x, y))
    assert D.AUTOTUNE == D.experimental.AUTOTUNE == -1
    out = list(ds.map(lambda a, b: (a * 2, b), num_parallel_calls=D.AUTOTUNE).batch(8).prefetch(D.AUTOTUNE))
    assert [o[0].shape for o in out] == [(8, 2), (8, 2), (4, 2)] and out[0][0][1, 0] == 4.0
    kept = list(ds.filter(lambda a, b: b == 1))
    assert [int(b) for _, b in kept] == [1] * 7 and float(kept[0][0][0]) == 2.0
    pairs = list(D.Dataset.zip((ds, D.Dataset.range(5))))
    assert len(pairs) == 5 and int(pairs[4][1]) == 4 and float(pairs[4][0][0][0]) == 8.0
    assert len(list(D.Dataset.zip(D.Dataset.range(3), D.Dataset.range(7)))) == 3
    assert int(D.Dataset.range(5).reduce(np.int64(0), lambda s, e: s + e)) == 10

    s = T.MirroredStrategy(devices=["cpu", "cpu"], cross_device_ops=T.NcclAllReduce())
    assert s.num_replicas_in_sync == 2 and s.cross_device_ops.name == "local"  # CPU replicas: no NVLink to use
    s2 = T.MirroredStrategy(devices=["cpu", "cpu"], cross_device_ops=T.HierarchicalCopyAllReduce())
    assert s2.cross_device_ops.name == "hierarchical_copy"
    g = [torch.ones(4), 2 * torch.ones(4)]
    T.ReductionToOneDevice().allreduce_grads(g)
    assert g[0].tolist() == g[1].tolist() == [3.0] * 4
    opts = T.experimental.CommunicationOptions(implementation=T.experimental.CommunicationImplementation.RING)
    single = T.MultiWorkerMirroredStrategy(communication_options=opts, devices=["cpu"], tf_config={})
    assert single.communication is T.CollectiveCommunication.RING and single.num_replicas_in_sync == 1
