"""model.save() / keras.models.load_model(): architecture + compile arguments + variables in one directory (the chief's
"checkpoint models" of the reference's README.md:51, as a whole model rather than weights only)."""
import json
import os

import numpy as np
import pytest

import tfdl
from tfdl import keras


def _data(n=96, seed=0):
    rng = np.random.default_rng(seed)
    return rng.random((n, 28, 28, 1), dtype=np.float32), rng.integers(0, 10, size=n)


def test_layer_configs_round_trip():
    L = keras.layers
    for layer in (L.Conv2D(8, (3, 5), strides=2, padding="same", activation="relu", use_bias=False, name="c"),
                  L.MaxPooling2D(3, 2, padding="same"), L.Dense(7, activation="relu"), L.Dropout(0.25), L.ReLU(),
                  L.Activation("relu"), L.BatchNormalization(momentum=0.9, epsilon=1e-4), L.LayerNormalization(epsilon=1e-6),
                  L.Embedding(11, 4), L.Flatten(), L.GlobalAveragePooling2D(), L.InputLayer((4, 4, 1))):
        cfg = layer.get_config()
        json.dumps(cfg)  # plain JSON values only
        twin = type(layer).from_config(dict(cfg))
        assert type(twin) is type(layer) and twin.get_config() == cfg, (type(layer).__name__, cfg, twin.get_config())
    c = L.Conv2D(8, (3, 5), strides=2, padding="same", activation="relu", use_bias=False, name="c")
    assert c.get_config() == {"filters": 8, "kernel_size": [3, 5], "strides": 2, "padding": "same", "activation": "relu",
                              "use_bias": False, "name": "c"}
    assert L.ReLU(name="r").get_config() == {"name": "r"}  # the outermost constructor's arguments, not Activation's


def test_save_and_load_sequential_with_optimizer_state(tmp_path):
    x, y = _data()
    model = tfdl.models.build_and_compile_cnn_model(learning_rate=0.05, momentum=0.9)
    model.fit(x, y, batch_size=32, epochs=1, shuffle=False, verbose=0)
    path = str(tmp_path / "saved")
    model.save(path)
    assert sorted(os.listdir(path)) == ["model.json", "variables"]
    spec = json.load(open(os.path.join(path, "model.json")))
    assert spec["class_name"] == "Sequential" and [l["class_name"] for l in spec["config"]["layers"]] == [
        "Conv2D", "MaxPooling2D", "Conv2D", "MaxPooling2D", "Flatten", "Dense", "Dense"]
    with pytest.raises(FileExistsError):
        model.save(path, overwrite=False)

    twin = keras.models.load_model(path)
    assert [v.name for v in twin.weights] == [v.name for v in model.weights]
    for a, b in zip(model.get_weights(), twin.get_weights()):
        np.testing.assert_array_equal(a, b)
    assert twin.optimizer.get_config() == model.optimizer.get_config()
    assert twin.optimizer.iterations == model.optimizer.iterations == 3
    assert twin.loss.from_logits and twin.metrics_names == ["loss", "sparse_categorical_accuracy"]
    # the momentum slots came along: the next steps of the original and of the restored model are the same steps
    h1 = model.fit(x, y, batch_size=32, epochs=1, shuffle=False, verbose=0)
    h2 = twin.fit(x, y, batch_size=32, epochs=1, shuffle=False, verbose=0)
    np.testing.assert_allclose(h1.history["loss"], h2.history["loss"], rtol=1e-6)
    for a, b in zip(model.get_weights(), twin.get_weights()):
        np.testing.assert_allclose(a, b, rtol=1e-5, atol=1e-7)

    plain = keras.models.load_model(path, compile=False)
    assert plain.optimizer is None
    np.testing.assert_allclose(plain.predict(x[:8]), keras.models.load_model(path).predict(x[:8]), rtol=1e-6)


def test_json_round_trip_and_custom_layer(tmp_path):
    class Scale(keras.layers.Layer):
        def __init__(self, factor: float = 2.0, **kw):
            super().__init__(**kw)
            self.factor = factor

        def call(self, x, ctx):
            return x * self.factor

    model = keras.Sequential([keras.layers.Flatten(input_shape=(4, 4, 1)), Scale(0.5), keras.layers.Dense(3)])
    model.compile(optimizer=keras.optimizers.Adam(1e-3, beta_2=0.99), loss="mse")
    text = model.to_json()
    with pytest.raises(ValueError, match="custom_objects"):
        keras.models.model_from_json(text)
    twin = keras.models.model_from_json(text, custom_objects={"Scale": Scale})
    assert isinstance(twin.layers[1], Scale) and twin.layers[1].factor == 0.5
    assert twin.optimizer.get_config() == model.optimizer.get_config()
    assert twin.count_params() == model.count_params() == 16 * 3 + 3
    path = str(tmp_path / "m")
    model.save(path)
    back = keras.models.load_model(path, custom_objects={"Scale": Scale})
    x = np.random.default_rng(1).random((5, 4, 4, 1), dtype=np.float32)
    np.testing.assert_allclose(back.predict(x), model.predict(x), rtol=1e-6)
    with pytest.raises(FileNotFoundError):
        keras.models.load_model(str(tmp_path / "nothing"))


def test_subclassed_model_families_save_and_load(tmp_path):
    """ResNet-50 and the BERT MLM model are subclassed models: their constructor arguments are the config."""
    res = tfdl.models.resnet50.ResNet50(classes=5, input_shape=(32, 32, 3))
    res.compile(optimizer=keras.optimizers.SGD(0.1, momentum=0.9), loss=keras.losses.SparseCategoricalCrossentropy(from_logits=True))
    path = str(tmp_path / "resnet")
    res.save(path)
    spec = json.load(open(os.path.join(path, "model.json")))
    assert spec["class_name"] == "ResNet50" and spec["config"]["classes"] == 5 and spec["config"]["input_shape"] == [32, 32, 3]
    twin = keras.models.load_model(path)
    assert type(twin).__name__ == "ResNet50" and twin.count_params() == res.count_params()
    for a, b in zip(res.get_weights()[:6] + res.get_weights()[-2:], twin.get_weights()[:6] + twin.get_weights()[-2:]):
        np.testing.assert_array_equal(a, b)


def _w_save(rank, addrs, outdir):
    os.environ["TF_CONFIG"] = json.dumps({"cluster": {"worker": addrs}, "task": {"type": "worker", "index": rank}})
    import tfdl
    from tfdl.models import build_and_compile_cnn_model

    s = tfdl.distribute.experimental.MultiWorkerMirroredStrategy(timeout_s=30)
    with s.scope():
        m = build_and_compile_cnn_model()
    x, y = _data(64)
    m.fit(x, y, batch_size=32, epochs=1, verbose=0, shuffle=False)
    mine = os.path.join(outdir, f"rank{rank}")
    m.save(mine)  # every worker calls it
    s.barrier()
    out = sorted(os.listdir(mine)) if os.path.isdir(mine) else []
    if rank == 0:
        with s.scope():
            twin = tfdl.keras.models.load_model(mine)
        same = all(np.array_equal(a, b) for a, b in zip(m.get_weights(), twin.get_weights()))
    else:
        same = None
    s.close()
    return out, same


def test_only_the_chief_writes_the_saved_model(tmp_path):
    from _mp import run_workers

    res = run_workers(_w_save, 2, args=(str(tmp_path),), timeout=120)
    assert res[0][0] == "ok" and res[1][0] == "ok", res
    assert res[0][1] == (["model.json", "variables"], True)
    assert res[1][1] == ([], None)  # README.md:51: the chief does the saving


def test_sidecar_evaluator_follows_the_chiefs_checkpoints(tmp_path):
    """The `evaluator` role of TF_CONFIG (README.md:38-44,57 of the reference): outside the training group, it evaluates
    each checkpoint the chief writes."""
    import threading

    from tfdl.cluster import TFConfigClusterResolver

    cfg = {"cluster": {"worker": ["127.0.0.1:1", "127.0.0.1:2"], "evaluator": ["127.0.0.1:3"]},
           "task": {"type": "evaluator", "index": 0}}
    res = TFConfigClusterResolver.from_env(cfg)
    assert res.task_type == "evaluator" and not res.is_training_task
    with pytest.raises(ValueError, match="not a training replica"):
        tfdl.distribute.MultiWorkerMirroredStrategy(tf_config=cfg, devices=["cpu"])

    x, y = _data(96)
    ckpt = str(tmp_path / "ckpt")
    trainer = tfdl.models.build_and_compile_cnn_model(learning_rate=0.05)
    side = tfdl.models.build_and_compile_cnn_model(learning_rate=0.05)  # the evaluator's own copy, default strategy
    seen = []
    ev = keras.utils.SidecarEvaluator(side, tfdl.data.Dataset.from_tensor_slices((x, y)).batch(32), ckpt, max_evaluations=2,
                                      callbacks=[keras.callbacks.LambdaCallback(on_epoch_end=lambda e, logs: seen.append(dict(logs)))],
                                      log_dir=str(tmp_path / "logs"), poll_interval_s=0.05, timeout_s=60)
    assert ev.evaluate_once() is None  # nothing written yet
    t = threading.Thread(target=ev.start)
    t.start()
    for epoch in range(2):  # the chief: one checkpoint per epoch
        trainer.fit(x, y, batch_size=32, epochs=1, shuffle=False, verbose=0,
                    callbacks=[keras.callbacks.ModelCheckpoint(os.path.join(ckpt, "ck-%d" % epoch))])
        want = trainer.evaluate(x, y, batch_size=32, return_dict=True)
        deadline = __import__("time").time() + 60
        while len(ev.history) <= epoch and __import__("time").time() < deadline:
            __import__("time").sleep(0.05)
        assert len(ev.history) == epoch + 1, ev.history
        got = ev.history[epoch]
        assert got["checkpoint"].endswith("ck-%d" % epoch) and got["step"] == 3 * (epoch + 1)
        assert abs(got["loss"] - want["loss"]) < 1e-6
    t.join(timeout=60)
    assert not t.is_alive() and len(seen) == 2 and "loss" in seen[0]
    assert os.listdir(tmp_path / "logs" / "validation")
