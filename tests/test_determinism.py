"""tfdl.config: seeds and the op-determinism switch (the counterpart of tf.keras.utils.set_random_seed and
tf.config.experimental.enable_op_determinism for users of the reference's stack)."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_RUN = r"""
import sys, json, hashlib
sys.path.insert(0, %(root)r)
import numpy as np
import tfdl
tfdl.config.set_random_seed(1234)
if %(det)r:
    tfdl.config.enable_op_determinism()
strategy = tfdl.distribute.MirroredStrategy(%(devices)s)
splits = tfdl.datasets.load(name='mnist', as_supervised=True, synthetic_size=(2048, 64), data_dir='/nonexistent')
ds = splits['train'].map(lambda x, y: (tfdl.utils.cast(x, tfdl.utils.float32) / 255, y)).cache().shuffle(500).batch(64)
with strategy.scope():
    model = tfdl.models.build_and_compile_cnn_model(learning_rate=0.05)
hist = model.fit(ds, epochs=2, steps_per_epoch=6, verbose=0)
w = np.concatenate([np.asarray(v).reshape(-1) for v in model.get_weights()])
print('RESULT ' + json.dumps({'engine': model.engine_name, 'sha': hashlib.sha256(w.tobytes()).hexdigest(),
                             'loss': hist.history['loss'], 'det': tfdl.config.is_op_determinism_enabled()}))
"""


def _run(det: bool, devices: str, extra_env=None):
    import json

    env = {k: v for k, v in os.environ.items() if k not in ("TF_CONFIG", "RANK", "WORLD_SIZE", "LOCAL_RANK")}
    env.update(extra_env or {})
    r = subprocess.run([sys.executable, "-c", _RUN % {"root": ROOT, "det": det, "devices": devices}], stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-3000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1]
    return json.loads(line[len("RESULT "):])


def test_seeded_cpu_runs_are_identical():
    a = _run(True, "devices=['cpu']", {"CUDA_VISIBLE_DEVICES": ""})
    b = _run(True, "devices=['cpu']", {"CUDA_VISIBLE_DEVICES": ""})
    assert a["det"] and a["engine"] == "TorchEngine"
    assert a["sha"] == b["sha"] and a["loss"] == b["loss"]


def test_switch_state_and_collective_choice():
    import torch

    import tfdl
    from tfdl.ops import allreduce as ar
    from tfdl.ops.structs import AR_MULTIMEM, AR_TWOSHOT

    assert not tfdl.config.is_op_determinism_enabled()
    before = (torch.are_deterministic_algorithms_enabled(), torch.backends.cudnn.deterministic, torch.backends.cudnn.benchmark)
    big = 64 << 20
    assert ar.pick_algo(big, 8, True) == AR_MULTIMEM
    tfdl.config.enable_op_determinism()
    try:
        assert tfdl.config.is_op_determinism_enabled() and torch.are_deterministic_algorithms_enabled()
        assert torch.backends.cudnn.deterministic and not torch.backends.cudnn.benchmark
        assert ar.pick_algo(big, 8, True) == AR_TWOSHOT  # no in-switch reduction: its order is not specified
        assert not ar.multimem_ok(8, True)
        for nbytes in (1 << 10, 1 << 16, 1 << 20, 4 << 20, big):
            assert ar.pick_algo(nbytes, 8, True) != AR_MULTIMEM
    finally:
        tfdl.config.disable_op_determinism()
    assert not tfdl.config.is_op_determinism_enabled()
    assert (torch.are_deterministic_algorithms_enabled(), torch.backends.cudnn.deterministic,
            torch.backends.cudnn.benchmark) == before
    assert ar.pick_algo(big, 8, True) == AR_MULTIMEM


def test_set_random_seed_fixes_initial_weights_and_shuffle_order():
    import tfdl

    def draw():
        tfdl.config.set_random_seed(7)
        m = tfdl.models.build_and_compile_cnn_model()
        m.build((28, 28, 1))
        ds = tfdl.data.Dataset.from_tensor_slices(np.arange(100)).shuffle(30).batch(10)
        return np.concatenate([np.asarray(v).reshape(-1) for v in m.get_weights()]), np.concatenate([np.asarray(b) for b in ds])

    try:
        (w1, o1), (w2, o2) = draw(), draw()
        assert np.array_equal(w1, w2) and np.array_equal(o1, o2)
        assert not np.array_equal(o1, np.arange(100))
    finally:
        tfdl.config.set_random_seed(None)


@pytest.mark.gpu
def test_deterministic_gpu_runs_are_bitwise_identical():
    """Two fresh processes, same seed, switch on: the same weights to the last bit, on the portable engine with library
    kernels in their deterministic modes and this framework's fused update."""
    a = _run(True, "devices=['gpu:0']")
    b = _run(True, "devices=['gpu:0']")
    assert a["det"] and a["engine"] == "TorchEngine"
    assert a["sha"] == b["sha"] and a["loss"] == b["loss"]
    fast = _run(False, "devices=['gpu:0']")
    assert fast["engine"] == "NativeCnnEngine" and not fast["det"]  # the default stays the native step program
