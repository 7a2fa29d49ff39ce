"""One TF_CONFIG worker driving one GPU: the reference script's own launch form (README.md:156-162, several cluster
nodes on one machine, README.md:61) with MultiWorkerMirroredStrategy(AUTO) -> fused NVLink all-reduce."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfdl  # noqa: E402


def main() -> None:
    cfg = json.loads(os.environ["TF_CONFIG"])
    idx = cfg["task"]["index"]
    strategy = tfdl.distribute.experimental.MultiWorkerMirroredStrategy(
        tfdl.distribute.experimental.CollectiveCommunication.AUTO, devices=[f"gpu:{idx}"])
    n = strategy.num_replicas_in_sync
    tfdl.keras.layers.set_init_seed(100 + idx)  # different local inits: the chief's values must win
    rng = np.random.default_rng(0)
    x = rng.random((64 * n * 5, 28, 28, 1), dtype=np.float32)
    y = rng.integers(0, 10, size=64 * n * 5).astype(np.int32)
    ds = tfdl.data.Dataset.from_tensor_slices((x, y)).batch(64 * n)
    options = tfdl.data.Options()
    options.experimental_distribute.auto_shard_policy = tfdl.data.experimental.AutoShardPolicy.OFF
    ds = ds.with_options(options)
    with strategy.scope():
        model = tfdl.models.build_and_compile_cnn_model(learning_rate=0.05)
    h = model.fit(x=ds, epochs=2, steps_per_epoch=5, verbose=0)
    w = np.concatenate([a.reshape(-1) for a in model.get_weights()])
    print("TFCFG_REPORT " + json.dumps({"index": idx, "replicas": n, "engine": model.engine_name,
                                        "ops": type(strategy.cross_device_ops).__name__, "loss": h.history["loss"],
                                        "wsum": float(np.float64(w).sum()), "wabs": float(np.abs(np.float64(w)).sum())}),
          flush=True)
    strategy.close()


if __name__ == "__main__":
    main()
