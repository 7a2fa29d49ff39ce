"""MultiWorkerMirroredStrategy on localhost processes (README.md:61 trick): TF_CONFIG per process,
TCP ring all-reduce, weights bit-identical on all workers and equal to a single-process run."""
import json
import os

import numpy as np
import pytest

from _mp import run_workers


def _train(rank, addrs, policy, comm, kill=None):
    import tfdl
    from tfdl import keras
    from tfdl.data import AutoShardPolicy, Dataset, Options
    from tfdl.models import build_and_compile_cnn_model

    os.environ["TF_CONFIG"] = json.dumps({"cluster": {"worker": addrs}, "task": {"type": "worker", "index": rank}})
    keras.layers.set_init_seed(100 + rank)  # different local inits: the chief's values must win
    strategy = tfdl.distribute.experimental.MultiWorkerMirroredStrategy(
        getattr(tfdl.distribute.experimental.CollectiveCommunication, comm), timeout_s=30, heartbeat_timeout_s=2)
    rng = np.random.default_rng(0)
    x, y = rng.random((256, 28, 28, 1), dtype=np.float32), rng.integers(0, 10, size=256)
    ds = Dataset.from_tensor_slices((x, y)).batch(64)  # GLOBAL batch
    o = Options()
    o.experimental_distribute.auto_shard_policy = getattr(AutoShardPolicy, policy)
    ds = ds.with_options(o)
    with strategy.scope():
        model = build_and_compile_cnn_model(learning_rate=0.05)
    cbs = [keras.callbacks.FaultInjection(kill)] if kill else []
    try:
        h = model.fit(ds, epochs=2, verbose=0, callbacks=cbs)
    except Exception as e:  # noqa: BLE001
        return {"error": f"{type(e).__name__}: {e}"}
    out = {"weights": model.get_weights(), "loss": h.history["loss"], "chief": strategy.is_chief,
           "replicas": strategy.num_replicas_in_sync, "comm": strategy.cross_device_ops.name, "steps": model._steps_done}
    strategy.close()
    return out


def _single(policy_data: bool):
    import tfdl
    from tfdl import keras
    from tfdl.data import Dataset
    from tfdl.models import build_and_compile_cnn_model

    keras.clear_session()
    keras.layers.set_init_seed(100)
    rng = np.random.default_rng(0)
    x, y = rng.random((256, 28, 28, 1), dtype=np.float32), rng.integers(0, 10, size=256)
    m = build_and_compile_cnn_model(learning_rate=0.05)
    h = m.fit(Dataset.from_tensor_slices((x, y)).batch(64), epochs=2, verbose=0)
    return m.get_weights(), h.history["loss"]


@pytest.mark.parametrize("comm", ["AUTO", "RING"])
def test_two_workers_data_sharded_equals_single_process(comm):
    res = run_workers(_train, 2, args=("DATA", comm), timeout=180)
    assert all(v[0] == "ok" and "error" not in v[1] for v in res.values()), res
    r0, r1 = res[0][1], res[1][1]
    assert r0["chief"] and not r1["chief"] and r0["replicas"] == 2 and r0["comm"] == "ring"
    for a, b in zip(r0["weights"], r1["weights"]):
        assert np.array_equal(a, b)  # bit-identical mirrored variables
    # DATA sharding: worker k takes elements k, k+2, ... so each global step covers the same 64 samples
    # as the single-process run on batch 64 -> same trajectory up to fp summation order
    ws, loss = _single(True)
    for a, b in zip(r0["weights"], ws):
        np.testing.assert_allclose(a, b, rtol=5e-4, atol=5e-6)
    np.testing.assert_allclose(r0["loss"], loss, rtol=1e-4)


def test_policy_off_every_worker_iterates_everything():
    res = run_workers(_train, 2, args=("OFF", "RING"), timeout=180)
    assert all(v[0] == "ok" and "error" not in v[1] for v in res.values()), res
    # 256 samples / (64 global / 2 workers = 32 per worker-step) = 8 steps per epoch on every worker
    assert res[0][1]["steps"] == 16 and res[1][1]["steps"] == 16
    for a, b in zip(res[0][1]["weights"], res[1][1]["weights"]):
        assert np.array_equal(a, b)


def test_four_workers_stay_identical_and_bad_split_is_rejected():
    res = run_workers(_train, 3, args=("OFF", "AUTO"), timeout=240)
    assert all("not divisible by 3 replicas" in v[1].get("error", "") for v in res.values()), res
    res = run_workers(_train, 4, args=("OFF", "AUTO"), timeout=240)
    assert all(v[0] == "ok" and "error" not in v[1] for v in res.values()), res
    for r in (1, 2, 3):
        for a, b in zip(res[0][1]["weights"], res[r][1]["weights"]):
            assert np.array_equal(a, b)


def test_killed_worker_surfaces_as_error_on_survivors():
    res = run_workers(_train, 2, args=("OFF", "RING", "kill:1:3"), timeout=120, expect_fail=(1,))
    status, val = res[0]
    assert status == "ok" and "error" in val, res
    assert "rank 1" in val["error"], val["error"]


def _roles(rank, addrs):
    import tfdl

    cfg = {"cluster": {"worker": addrs, "ps": ["127.0.0.1:1"]}, "task": {"type": "ps", "index": 0}}
    try:
        tfdl.distribute.experimental.MultiWorkerMirroredStrategy(tf_config=cfg)
    except ValueError as e:
        return str(e)
    return "no error"


def test_ps_role_is_parsed_but_rejected():
    res = run_workers(_roles, 1)
    assert "not a training replica" in res[0][1]


def _topology(rank, addrs, comm):
    """Workers that report different hosts: the reference's own two-machine deployment (tf_dist_example.py:7)."""
    import tfdl

    os.environ["TFDL_HOST_ID"] = f"machine-{rank}"
    os.environ["TF_CONFIG"] = json.dumps({"cluster": {"worker": addrs}, "task": {"type": "worker", "index": rank}})
    try:
        s = tfdl.distribute.experimental.MultiWorkerMirroredStrategy(
            getattr(tfdl.distribute.experimental.CollectiveCommunication, comm), timeout_s=30, heartbeat_timeout_s=2)
    except ValueError as e:
        return "ValueError: " + str(e)
    name = s.cross_device_ops.name
    s.close()
    return name


def test_auto_across_hosts_takes_the_ring_and_nccl_is_refused():
    res = run_workers(_topology, 2, args=("AUTO",), timeout=60)
    assert res[0][1] == "ring" and res[1][1] == "ring", res
    res = run_workers(_topology, 2, args=("NCCL",), timeout=60)
    assert all("different hosts" in v[1] for v in res.values()), res


def _train_with_validation(rank, addrs):
    import tfdl
    from tfdl import keras
    from tfdl.models import build_and_compile_cnn_model

    os.environ["TF_CONFIG"] = json.dumps({"cluster": {"chief": addrs[:1], "worker": addrs[1:]},
                                          "task": {"type": "chief" if rank == 0 else "worker", "index": max(0, rank - 1)}})
    strategy = tfdl.distribute.experimental.MultiWorkerMirroredStrategy(timeout_s=30, heartbeat_timeout_s=5)
    rng = np.random.default_rng(0)
    x, y = rng.random((192, 28, 28, 1), dtype=np.float32), rng.integers(0, 10, size=192)
    xv, yv = rng.random((60, 28, 28, 1), dtype=np.float32), rng.integers(0, 10, size=60)
    with strategy.scope():
        model = build_and_compile_cnn_model(learning_rate=0.05)
    cb = keras.callbacks.EarlyStopping(monitor="val_loss", patience=0, min_delta=100.0)  # stops after the second epoch
    h = model.fit(x, y, batch_size=48, epochs=5, shuffle=False, verbose=0, validation_data=(xv, yv), callbacks=[cb])
    alone = model.evaluate(xv, yv, batch_size=48, return_dict=True)
    out = {"history": h.history, "alone": alone, "weights": model.get_weights(), "chief": strategy.is_chief}
    strategy.close()
    return out


def test_three_tasks_with_a_chief_validation_and_early_stopping():
    """chief + 2 workers (README.md:38-51 of the reference), arrays in, validation arrays, a callback that stops training
    from the validation loss: every task sees the same logs, takes the same decision and ends with the same weights."""
    res = run_workers(_train_with_validation, 3, timeout=240)
    assert all(v[0] == "ok" for v in res.values()), res
    outs = [res[r][1] for r in range(3)]
    assert [o["chief"] for o in outs] == [True, False, False]
    for o in outs[1:]:
        assert o["history"] == outs[0]["history"]
        for a, b in zip(o["weights"], outs[0]["weights"]):
            np.testing.assert_array_equal(a, b)
    h = outs[0]["history"]
    assert len(h["loss"]) == 2 and len(h["val_loss"]) == 2  # stopped early, everywhere at the same epoch
    assert abs(h["val_loss"][-1] - outs[0]["alone"]["loss"]) < 1e-5
