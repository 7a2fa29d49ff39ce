"""Keras-like front end on CPU: shapes, parameter counts, fp64 oracle for the step math, replica
equivalence (N replicas at global batch G == 1 replica at batch G), fit semantics."""
import numpy as np
import pytest
import torch

import tfdl
from tfdl import keras
from tfdl.data import Dataset
from tfdl.models import build_and_compile_cnn_model, build_cnn_model


def _fresh(seed=0):
    keras.clear_session()
    keras.layers.set_init_seed(seed)


def _data(n=256, seed=0):
    rng = np.random.default_rng(seed)
    return rng.random((n, 28, 28, 1), dtype=np.float32), rng.integers(0, 10, size=n)


def test_reference_model_shapes_names_and_count():
    _fresh()
    m = build_cnn_model()
    want = [("conv2d/kernel:0", (3, 3, 1, 32)), ("conv2d/bias:0", (32,)), ("conv2d_1/kernel:0", (3, 3, 32, 64)),
            ("conv2d_1/bias:0", (64,)), ("dense/kernel:0", (1600, 128)), ("dense/bias:0", (128,)),
            ("dense_1/kernel:0", (128, 10)), ("dense_1/bias:0", (10,))]
    assert [(w.name, w.shape) for w in m.weights] == want
    assert m.count_params() == 225034
    assert [l.output_shape for l in m.layers] == [(26, 26, 32), (13, 13, 32), (11, 11, 64), (5, 5, 64), (1600,), (128,), (10,)]
    assert m.output_shape == (None, 10)
    # Glorot-uniform kernels within their limit, zero biases
    k = m.get_weights()[4]
    assert np.abs(k).max() <= np.sqrt(6 / (1600 + 128)) + 1e-6 and np.all(m.get_weights()[5] == 0)


def _numpy_forward_backward(ws, x, y):
    """fp64 restatement of the model math (conv, pool, dense, softmax-xent) and its gradients."""
    k1, b1, k2, b2, w3, b3, w4, b4 = [w.astype(np.float64) for w in ws]
    x = x.astype(np.float64)

    def conv(a, k, b):
        n, h, w, c = a.shape
        kh, kw, _, co = k.shape
        cols = np.stack([a[:, i:i + h - kh + 1, j:j + w - kw + 1, :] for i in range(kh) for j in range(kw)], axis=3)
        return np.einsum("nhwtc,tco->nhwo", cols, k.reshape(kh * kw, c, co)) + b, cols

    def pool(a):
        n, h, w, c = a.shape
        a = a[:, :h // 2 * 2, :w // 2 * 2, :].reshape(n, h // 2, 2, w // 2, 2, c)
        return a.max(axis=(2, 4))

    z1, _ = conv(x, k1, b1)
    a1 = pool(np.maximum(z1, 0))
    z2, _ = conv(a1, k2, b2)
    a2 = pool(np.maximum(z2, 0))
    f = a2.reshape(len(x), -1)
    h = np.maximum(f @ w3 + b3, 0)
    logits = h @ w4 + b4
    logits -= logits.max(1, keepdims=True)
    logp = logits - np.log(np.exp(logits).sum(1, keepdims=True))
    loss = -logp[np.arange(len(y)), y].mean()
    dlog = np.exp(logp)
    dlog[np.arange(len(y)), y] -= 1
    dlog /= len(y)
    return loss, {"w4": h.T @ dlog, "b4": dlog.sum(0), "dh": dlog @ w4.T * (h > 0), "f": f}


def test_step_matches_fp64_numpy_oracle():
    _fresh(1)
    m = build_and_compile_cnn_model(learning_rate=0.1)
    x, y = _data(32, 1)
    ws = m.get_weights()
    loss_ref, g = _numpy_forward_backward(ws, x, y)
    logs = m.train_on_batch(x, y, return_dict=True)
    assert abs(logs["loss"] - loss_ref) < 1e-5
    new = m.get_weights()
    np.testing.assert_allclose((ws[6] - new[6]) / 0.1, g["w4"], rtol=1e-3, atol=1e-6)   # dense_1/kernel gradient
    np.testing.assert_allclose((ws[7] - new[7]) / 0.1, g["b4"], rtol=1e-3, atol=1e-6)
    np.testing.assert_allclose((ws[4] - new[4]) / 0.1, g["f"].T @ g["dh"], rtol=1e-3, atol=1e-6)  # dense/kernel (h,w,c flatten order)


@pytest.mark.parametrize("nrep", [2, 4])
def test_n_replicas_equal_one_replica_at_same_global_batch(nrep):
    x, y = _data(128, 2)
    results = []
    for devs in (["cpu:0"], ["cpu:%d" % i for i in range(nrep)]):
        _fresh(5)
        s = tfdl.distribute.MirroredStrategy(devices=devs)
        with s.scope():
            m = build_and_compile_cnn_model(learning_rate=0.05, momentum=0.9)
        ds = Dataset.from_tensor_slices((x, y)).batch(64)
        h = m.fit(ds, epochs=2, verbose=0)
        # all replica copies identical after every update (mirrored variables)
        for w in m.weights:
            for r in range(1, s.num_local_replicas):
                assert torch.equal(w.value(0), w.value(r)), w.name
        results.append((m.get_weights(), h.history["loss"]))
    for a, b in zip(results[0][0], results[1][0]):
        np.testing.assert_allclose(a, b, rtol=2e-4, atol=2e-6)
    np.testing.assert_allclose(results[0][1], results[1][1], rtol=1e-4)


def test_fit_carries_iterator_across_epochs_with_steps_per_epoch(capsys):
    _fresh()
    x, y = _data(64 * 5, 3)
    seen = []
    ds = Dataset.from_tensor_slices((x, y)).map(lambda a, b: (seen.append(int(b[0])) or a, b), vectorized=True).batch(64)
    m = build_and_compile_cnn_model()
    h = m.fit(ds, epochs=2, steps_per_epoch=2, verbose=0)
    assert len(h.history["loss"]) == 2 and m._steps_done == 4
    assert seen[:4] == [int(y[0]), int(y[64]), int(y[128]), int(y[192])]  # no restart between epochs
    # running dry: warns and stops (the reference relies on 200 steps < one pass, quirk 7)
    h = m.fit(ds, epochs=3, steps_per_epoch=4, verbose=1)
    assert "ran out of data" in capsys.readouterr().err and len(h.history["loss"]) == 2


def test_training_learns_and_evaluate_predict():
    _fresh(7)
    ds = tfdl.datasets.load("mnist", as_supervised=True, synthetic_size=(1024, 128))
    prep = lambda i, l: (tfdl.utils.cast(i, "float32") / 255, l)  # noqa: E731
    train, test = ds["train"].map(prep).cache().shuffle(1024, seed=0).batch(64), ds["test"].map(prep).batch(64)
    m = build_and_compile_cnn_model(learning_rate=0.05)
    h = m.fit(train, epochs=3, verbose=0, validation_data=test)
    assert h.history["loss"][-1] < 0.5 * h.history["loss"][0]
    assert h.history["val_sparse_categorical_accuracy"][-1] > 0.8
    ev = m.evaluate(test, return_dict=True)
    assert abs(ev["loss"] - h.history["val_loss"][-1]) < 1e-5
    p = m.predict(next(iter(test))[0])
    assert p.shape == (64, 10)


def test_compile_requires_same_scope_and_api_errors():
    _fresh()
    s1, s2 = tfdl.distribute.MirroredStrategy(devices=["cpu:0"]), tfdl.distribute.MirroredStrategy(devices=["cpu:0"])
    with s1.scope():
        m = build_cnn_model()
    with s2.scope():
        with pytest.raises(RuntimeError, match="different strategy"):
            m.compile(optimizer="sgd", loss=keras.losses.SparseCategoricalCrossentropy(from_logits=True))
    with pytest.raises(RuntimeError, match="compile"):
        m.fit(*_data(32))
    with pytest.raises(ValueError):
        keras.layers.Conv2D(8, 3, activation="nope")


def test_callbacks_lr_schedule_and_early_stopping():
    _fresh()
    x, y = _data(128)
    m = build_and_compile_cnn_model(learning_rate=0.1)
    lrs = []
    cb = [keras.callbacks.LearningRateScheduler(lambda e, lr: lr * 0.5),
          keras.callbacks.LambdaCallback(on_epoch_end=lambda e, logs: lrs.append(m.optimizer.learning_rate)),
          keras.callbacks.EarlyStopping(monitor="loss", patience=0, min_delta=10.0)]
    h = m.fit(x, y, batch_size=64, epochs=5, verbose=0, callbacks=cb)
    assert lrs[:2] == [0.05, 0.025] and len(h.history["loss"]) == 2  # stopped: no 10.0 improvement


def test_fault_injection_is_installed_from_the_environment(monkeypatch):
    """TFDL_FAULT alone arms the fault hook: no callback has to be passed to fit()."""
    import tfdl
    from tfdl import keras

    seen = []
    orig = keras.callbacks.FaultInjection.on_train_batch_end

    def spy(self, batch, logs=None):
        seen.append((self.spec, batch))
        return orig(self, batch, logs)

    monkeypatch.setattr(keras.callbacks.FaultInjection, "on_train_batch_end", spy)
    monkeypatch.setenv("TFDL_FAULT", "stall:7:1000:0.0")  # rank 7 does not exist here: armed but never fires
    m = keras.Sequential([keras.layers.Dense(4, input_shape=(3,))])
    m.compile(optimizer="sgd", loss="mse")
    x, y = np.random.rand(16, 3).astype(np.float32), np.random.rand(16, 4).astype(np.float32)
    m.fit(x, y, batch_size=8, epochs=1, verbose=0)
    assert seen and seen[0][0] == "stall:7:1000:0.0"


def test_relu_after_pool_on_cpu_is_exactly_relu_before_pool():
    """CPU workers run Conv2D(relu) -> MaxPooling2D as conv -> pool -> relu (keras/engine.py:Sequential._forward). The two
    orders are the same function with the same gradients - also for windows that are all negative or hold ties."""
    import torch

    import tfdl
    from tfdl.keras import layers as L
    from tfdl.keras.layers import CallContext

    tfdl.config.set_random_seed(3)
    try:
        model = tfdl.models.build_and_compile_cnn_model()
        model.build((28, 28, 1))
        model._finalize()
    finally:
        tfdl.config.set_random_seed(None)
    rng = np.random.default_rng(0)
    x = rng.standard_normal((8, 28, 28, 1)).astype(np.float32)
    x[0] = -1.0       # whole images whose first-layer windows are all negative / all equal
    x[1] = 0.25
    x[2, :14] = 0.0
    pools = [l for l in model._stack if type(l) is L.MaxPooling2D]
    assert len(pools) == 2

    class Plain(L.MaxPooling2D):  # a subclass is not fused: the layer-by-layer order
        pass

    def run(fused: bool):
        for p in pools:
            p.__class__ = L.MaxPooling2D if fused else Plain
        leaves = {id(v): v.value(0).detach().clone().requires_grad_(True) for v in model.trainable_variables}
        ctx = CallContext(0, True, lambda v: leaves[id(v)], torch.float32)
        out = model._forward(torch.from_numpy(x), ctx)
        out.square().sum().backward()
        return out.detach(), [leaves[id(v)].grad for v in model.trainable_variables]

    try:
        (o1, g1), (o2, g2) = run(True), run(False)
    finally:
        for p in pools:
            p.__class__ = L.MaxPooling2D
    assert torch.equal(o1, o2)
    for a, b in zip(g1, g2):
        torch.testing.assert_close(a, b, rtol=1e-6, atol=1e-7)  # same terms; the library may add them in another order


def test_validation_data_arrays_and_validation_split():
    """fit(validation_data=(x_val, y_val)) - the arrays form the reference's README shows (README.md:121-129 builds
    train and val arrays) - and fit(validation_split=...): val_ metrics every epoch, the held-out tail never trained on."""
    import tfdl

    rng = np.random.default_rng(0)
    x, y = rng.random((96, 28, 28, 1), dtype=np.float32), rng.integers(0, 10, size=96)
    m = tfdl.models.build_and_compile_cnn_model(learning_rate=0.01)
    h = m.fit(x[:64], y[:64], batch_size=32, epochs=2, verbose=0, validation_data=(x[64:], y[64:]))
    assert set(h.history) == {"loss", "sparse_categorical_accuracy", "val_loss", "val_sparse_categorical_accuracy"}
    assert len(h.history["val_loss"]) == 2
    want = m.evaluate(x[64:], y[64:], batch_size=32, return_dict=True)["loss"]
    assert abs(h.history["val_loss"][-1] - want) < 1e-5  # validation ran on exactly those arrays, after the last epoch
    ds_val = tfdl.data.Dataset.from_tensor_slices((x[64:], y[64:])).batch(16)
    h2 = m.fit(x[:64], y[:64], batch_size=32, epochs=1, verbose=0, validation_data=ds_val)
    assert "val_loss" in h2.history

    m2 = tfdl.models.build_and_compile_cnn_model(learning_rate=0.01)
    h3 = m2.fit(x, y, batch_size=24, epochs=1, verbose=0, validation_split=0.25, shuffle=False)
    assert m2._steps_done == 3  # 72 training rows / 24, the last 24 rows held out
    want = m2.evaluate(x[72:], y[72:], batch_size=24, return_dict=True)["loss"]
    assert abs(h3.history["val_loss"][-1] - want) < 1e-5
    with pytest.raises(ValueError, match="validation_split"):
        m2.fit(ds_val, epochs=1, verbose=0, validation_split=0.5)
