"""MirroredStrategy() in ONE plain python process over all visible GPUs (the form /root/reference/tf_dist_example.py:13
names): must select the native multi-GPU engine, keep the replicas bit-identical and match a single-GPU run on the same
global batches. Launched by tests/test_multigpu.py::test_single_process_mirrored_strategy."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfdl  # noqa: E402
from tfdl import keras  # noqa: E402
from tfdl.models import build_and_compile_cnn_model  # noqa: E402


def main() -> None:
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        os.environ.pop(k, None)
    n = torch.cuda.device_count()
    strategy = tfdl.distribute.MirroredStrategy()
    assert strategy.num_replicas_in_sync == n and strategy.num_workers == 1, strategy
    keras.clear_session()
    keras.layers.set_init_seed(7)
    with strategy.scope():
        model = build_and_compile_cnn_model(learning_rate=0.05, momentum=0.9)
    rng = np.random.default_rng(0)
    G = 64 * n
    x, y = rng.random((G * 6, 28, 28, 1), dtype=np.float32), rng.integers(0, 10, size=G * 6)
    ds = tfdl.data.Dataset.from_tensor_slices((x, y)).shuffle(1000, seed=3).batch(G)
    w_init = model.get_weights()
    h = model.fit(ds, epochs=2, verbose=0)
    rep = {"gpus": n, "engine": model.engine_name, "ops": type(strategy.cross_device_ops).__name__,
           "algo": model._engine.algo_name, "loss": h.history["loss"]}
    # replicas bit-identical
    for v in model.weights:
        vals = [t.detach().cpu() for t in v.values]
        for t in vals[1:]:
            assert torch.equal(vals[0], t), f"replicas diverged in {v.name}"
    rep["replicas_identical"] = True
    # same run on one GPU with the same global batches
    keras.clear_session()
    keras.layers.set_init_seed(7)
    solo = tfdl.distribute.MirroredStrategy(devices=["gpu:0"])
    with solo.scope():
        ref = build_and_compile_cnn_model(learning_rate=0.05, momentum=0.9)
    ds2 = tfdl.data.Dataset.from_tensor_slices((x, y)).shuffle(1000, seed=3).batch(G)
    hr = ref.fit(ds2, epochs=2, verbose=0)
    np.testing.assert_allclose(h.history["loss"], hr.history["loss"], rtol=2e-2)
    worst = 0.0
    for a, b, w0 in zip(model.get_weights(), ref.get_weights(), w_init):
        ua, ub = (a - w0).astype(np.float64), (b - w0).astype(np.float64)
        worst = max(worst, float(np.linalg.norm(ua - ub) / (np.linalg.norm(ub) + 1e-30)))
    assert worst < 0.2, worst
    rep["update_rel_diff_vs_single_gpu"] = worst
    # a checkpoint sees complete momentum slots on replica 0 (sliced buckets keep them distributed between steps)
    strategy.cross_device_ops.sync_optimizer_slots()
    m0 = model.optimizer.slots("momentum")
    for t in m0[1:]:
        assert torch.equal(m0[0].cpu(), t.cpu()), "momentum slots differ between replicas after the gather"
    rep["momentum_gathered"] = True
    print("MIRRORED_REPORT " + json.dumps(rep), flush=True)


if __name__ == "__main__":
    main()
