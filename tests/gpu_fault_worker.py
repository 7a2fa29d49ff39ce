"""One TF_CONFIG worker on one GPU whose peer dies mid-run (launched by tests/test_multigpu.py::test_killed_gpu_worker).
Rank 1 exits at global step 3 (callbacks.FaultInjection); rank 0 must come back from its NVLink collective with an error
that names rank 1 within the heartbeat timeout - not spin into the kernels' trap limit, not hang."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tfdl  # noqa: E402


def main() -> None:
    cfg = json.loads(os.environ["TF_CONFIG"])
    idx = cfg["task"]["index"]
    strategy = tfdl.distribute.experimental.MultiWorkerMirroredStrategy(
        tfdl.distribute.experimental.CollectiveCommunication.AUTO, devices=[f"gpu:{idx}"], heartbeat_timeout_s=2.0)
    n = strategy.num_replicas_in_sync
    rng = np.random.default_rng(0)
    x = rng.random((64 * n * 8, 28, 28, 1), dtype=np.float32)
    y = rng.integers(0, 10, size=64 * n * 8).astype(np.int32)
    ds = tfdl.data.Dataset.from_tensor_slices((x, y)).batch(64 * n)
    options = tfdl.data.Options()
    options.experimental_distribute.auto_shard_policy = tfdl.data.experimental.AutoShardPolicy.OFF
    ds = ds.with_options(options)
    with strategy.scope():
        model = tfdl.models.build_and_compile_cnn_model(learning_rate=0.05)
    t0 = time.time()
    out = {"index": idx, "engine": None, "error": None}
    try:
        model.fit(x=ds, epochs=3, verbose=0, callbacks=[tfdl.keras.callbacks.FaultInjection("kill:1:3")])
        out["engine"] = model.engine_name
    except Exception as e:  # noqa: BLE001
        out["engine"] = model.engine_name
        out["error"] = f"{type(e).__name__}: {e}"
    out["seconds"] = time.time() - t0
    print("FAULT_REPORT " + json.dumps(out), flush=True)
    os._exit(0)  # the group is broken: no orderly shutdown to wait for


if __name__ == "__main__":
    main()
