"""The native step program (CUDA graph of sm_100a kernels) against the portable torch engine."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _data(n, seed=0):
    rng = np.random.default_rng(seed)
    return rng.random((n, 28, 28, 1), dtype=np.float32), rng.integers(0, 10, size=n)


def _model(engine, lr=0.05, momentum=0.0, seed=3):
    import tfdl
    from tfdl import keras
    from tfdl.models import build_and_compile_cnn_model

    keras.layers.set_init_seed(seed)
    keras.clear_session()
    s = tfdl.distribute.MirroredStrategy(devices=["gpu:0"])
    with s.scope():
        m = build_and_compile_cnn_model(learning_rate=lr, momentum=momentum)
    m._engine_pref = engine
    return m


@pytest.mark.parametrize("momentum", [0.0, 0.9])
def test_native_steps_track_torch_engine(momentum):
    x, y = _data(64 * 4)
    native, ref = _model("native", momentum=momentum), _model("torch", momentum=momentum)
    for a, b in zip(native.get_weights(), ref.get_weights()):
        np.testing.assert_array_equal(a, b)
    w0 = ref.get_weights()
    for i in range(4):
        xb, yb = x[i * 64:(i + 1) * 64], y[i * 64:(i + 1) * 64]
        ln = native.train_on_batch(xb, yb, return_dict=True)
        lr_ = ref.train_on_batch(xb, yb, return_dict=True)
        assert abs(ln["loss"] - lr_["loss"]) < 2e-2, (i, ln, lr_)
        assert abs(ln["sparse_categorical_accuracy"] - lr_["sparse_categorical_accuracy"]) <= 4 / 64
    assert native.engine_name == "NativeCnnEngine" and ref.engine_name == "TorchEngine"
    for name, a, b, w in zip([v.name for v in ref.weights], native.get_weights(), ref.get_weights(), w0):
        upd_ref = (b - w).astype(np.float64)
        diff = (a - w).astype(np.float64) - upd_ref
        rel_l2 = np.linalg.norm(diff) / (np.linalg.norm(upd_ref) + 1e-30)
        worst = np.abs(diff).max() / (np.abs(upd_ref).max() + 1e-30)
        # bf16 activations / activation gradients vs an fp32 reference: a few percent in the norm
        assert rel_l2 < 0.10 and worst < 0.30, (name, rel_l2, worst)


@pytest.mark.parametrize("batch", [32, 256])
def test_other_batch_sizes_track_torch_engine(batch):
    """Batch sizes outside the fused dense block's range (B > 128) run the unbucketed program; same updates."""
    x, y = _data(batch * 3, seed=5)
    native, ref = _model("native", momentum=0.9), _model("torch", momentum=0.9)
    w0 = ref.get_weights()
    for i in range(3):
        xb, yb = x[i * batch:(i + 1) * batch], y[i * batch:(i + 1) * batch]
        ln, lr_ = native.train_on_batch(xb, yb, return_dict=True), ref.train_on_batch(xb, yb, return_dict=True)
        assert abs(ln["loss"] - lr_["loss"]) < 2e-2, (i, ln, lr_)
    assert native.engine_name == "NativeCnnEngine"
    for name, a, b, w in zip([v.name for v in ref.weights], native.get_weights(), ref.get_weights(), w0):
        upd_ref = (b - w).astype(np.float64)
        rel_l2 = np.linalg.norm((a - w).astype(np.float64) - upd_ref) / (np.linalg.norm(upd_ref) + 1e-30)
        assert np.linalg.norm(upd_ref) > 0 and rel_l2 < 0.10, (name, rel_l2)


def test_fit_on_gpu_learns_and_uses_native_engine():
    import tfdl
    from tfdl import datasets

    ds = datasets.load(name="mnist", as_supervised=True, synthetic_size=(4096, 64))["train"]
    train = ds.map(lambda i, l: (tfdl.utils.cast(i, "float32") / 255, l)).cache().shuffle(4096, seed=1).batch(64).repeat()
    m = _model("auto", lr=0.05)
    h = m.fit(train, epochs=3, steps_per_epoch=40, verbose=0)
    assert m.engine_name == "NativeCnnEngine"
    assert h.history["loss"][-1] < h.history["loss"][0] * 0.7, h.history
    assert h.history["sparse_categorical_accuracy"][-1] > 0.6, h.history
    ev = m.evaluate(train, steps=5, return_dict=True)
    assert ev["loss"] < h.history["loss"][0]


def test_graph_and_eager_program_agree():
    import os

    x, y = _data(64 * 3, seed=5)
    outs = []
    for use_graph in ("1", "0"):
        os.environ["TFDL_USE_GRAPH"] = use_graph
        m = _model("native")
        for i in range(3):
            m.train_on_batch(x[i * 64:(i + 1) * 64], y[i * 64:(i + 1) * 64])
        outs.append(m.get_weights())
    os.environ.pop("TFDL_USE_GRAPH")
    for a, b in zip(*outs):
        # only the atomics' summation order may differ
        np.testing.assert_allclose(a, b, rtol=0, atol=2e-5)


def test_ragged_batch_falls_back_and_resyncs():
    x, y = _data(64 + 40, seed=9)
    m = _model("native")
    m.train_on_batch(x[:64], y[:64])
    logs = m.train_on_batch(x[64:], y[64:], return_dict=True)  # 40 is not a multiple of 16 -> torch step
    assert np.isfinite(logs["loss"])
    logs = m.train_on_batch(x[:64], y[:64], return_dict=True)
    assert np.isfinite(logs["loss"])
    eng = m._engine
    torch.testing.assert_close(eng.state.shadow.tensor.float(), eng.state.params.tensor.bfloat16().float())


def test_zero_staging_inputs_match_ring_path(monkeypatch):
    """Unshuffled from_tensor_slices batches are DMA'd from the page-locked dataset array itself; same training."""
    import tfdl

    x, y = _data(64 * 12, seed=9)
    y = y.astype(np.int32)
    results = []
    for zc in ("1", "0"):
        monkeypatch.setenv("TFDL_ZERO_COPY", zc)
        m = _model("native", lr=0.05)
        ds = tfdl.data.Dataset.from_tensor_slices((x, y)).batch(64)
        h = m.fit(ds, epochs=2, verbose=0)
        used = any(own is not None for own, _ in m._engine._pinned.values())
        assert used == (zc == "1")
        results.append((h.history["loss"], m.get_weights()))
        del m
    # same data, same kernels; split-K / RED accumulation order differs run to run, so not bit-identical
    np.testing.assert_allclose(results[0][0], results[1][0], rtol=2e-3)
    for a, b in zip(results[0][1], results[1][1]):
        np.testing.assert_allclose(a, b, rtol=0, atol=5e-3)


def test_torch_engine_whole_step_cuda_graph_matches_eager(monkeypatch):
    """TorchEngine captures the whole step (forward, backward, metrics, update) after three eager warm-up steps."""
    x, y = _data(64 * 7, seed=21)
    out = {}
    for mode in ("1", "0"):
        monkeypatch.setenv("TFDL_TORCH_GRAPH", mode)
        m = _model("torch", lr=0.05, momentum=0.9)
        losses = []
        for i in range(7):
            losses.append(m.train_on_batch(x[i * 64:(i + 1) * 64], y[i * 64:(i + 1) * 64], return_dict=True)["loss"])
        assert m.engine_name == "TorchEngine"
        assert bool(m._engine._graphs) == (mode == "1")
        out[mode] = (losses, m.get_weights())
    np.testing.assert_allclose(out["1"][0], out["0"][0], rtol=2e-3)
    for a, b in zip(out["1"][1], out["0"][1]):  # same kernels; library atomics make the two runs differ in the last bits
        np.testing.assert_allclose(a, b, rtol=0, atol=1e-3)


def test_native_program_with_adam_tracks_torch_engine():
    """Adam runs inside the native step program too (fused into the all-reduce + update kernels)."""
    import tfdl
    from tfdl import keras
    from tfdl.models import build_and_compile_cnn_model

    def model(engine):
        keras.layers.set_init_seed(3)
        keras.clear_session()
        s = tfdl.distribute.MirroredStrategy(devices=["gpu:0"])
        with s.scope():
            m = build_and_compile_cnn_model(learning_rate=1e-3)
            m.compile(loss=keras.losses.SparseCategoricalCrossentropy(from_logits=True),
                      optimizer=keras.optimizers.Adam(learning_rate=2e-3),
                      metrics=[keras.metrics.SparseCategoricalAccuracy()])
        m._engine_pref = engine
        return m

    x, y = _data(64 * 6, seed=8)
    native, ref = model("native"), model("torch")
    for i in range(6):
        xb, yb = x[i * 64:(i + 1) * 64], y[i * 64:(i + 1) * 64]
        ln, lr_ = native.train_on_batch(xb, yb, return_dict=True), ref.train_on_batch(xb, yb, return_dict=True)
        assert abs(ln["loss"] - lr_["loss"]) < 3e-2, (i, ln, lr_)
    assert native.engine_name == "NativeCnnEngine" and ref.engine_name == "TorchEngine"
    assert ln["loss"] < 2.30  # six Adam steps move the loss
