"""Chief-only TensorBundle checkpoints: Keras object-graph keys and layouts, an INDEPENDENT pure-Python
reader of the on-disk format, corruption detection, resume, TensorBoard events, chief-only writing."""
import json
import os
import struct

import numpy as np
import pytest

import tfdl
from tfdl import keras
from tfdl.checkpoint import keras_checkpoint_keys, latest_checkpoint
from tfdl.models import build_and_compile_cnn_model

from _mp import run_workers


# ------------------------------------------------------------------------------------------------
# independent reader: LevelDB table format + hand-rolled protobuf decoding, no code shared with the writer
# ------------------------------------------------------------------------------------------------
def _varint(buf, pos):
    out, shift = 0, 0
    while True:
        b = buf[pos]
        pos += 1
        out |= (b & 0x7F) << shift
        if not b & 0x80:
            return out, pos
        shift += 7


def _crc32c_py(data, crc=0):
    crc ^= 0xFFFFFFFF
    for byte in data:
        crc ^= byte
        for _ in range(8):
            crc = (crc >> 1) ^ (0x82F63B78 & -(crc & 1))
    return crc ^ 0xFFFFFFFF


def _unmask(m):
    rot = (m - 0xA282EAD8) & 0xFFFFFFFF
    return ((rot >> 17) | (rot << 15)) & 0xFFFFFFFF


def _read_block(f, off, size):
    raw = f[off:off + size]
    typ, crc = f[off + size], struct.unpack("<I", f[off + size + 1:off + size + 5])[0]
    assert typ == 0 and _unmask(crc) == _crc32c_py(f[off:off + size + 1])
    nrestart = struct.unpack("<I", raw[-4:])[0]
    end = len(raw) - 4 - 4 * nrestart
    pos, key, out = 0, b"", []
    while pos < end:
        shared, pos = _varint(raw, pos)
        non_shared, pos = _varint(raw, pos)
        vlen, pos = _varint(raw, pos)
        key = key[:shared] + raw[pos:pos + non_shared]
        pos += non_shared
        out.append((key, raw[pos:pos + vlen]))
        pos += vlen
    return out


def read_index(path):
    f = open(path, "rb").read()
    assert struct.unpack("<Q", f[-8:])[0] == 0xDB4775248B80FB57
    pos = len(f) - 48
    _, pos = _varint(f, pos)
    _, pos = _varint(f, pos)
    ioff, pos = _varint(f, pos)
    isize, pos = _varint(f, pos)
    entries = {}
    for _, handle in _read_block(f, ioff, isize):
        boff, p = _varint(handle, 0)
        bsize, p = _varint(handle, p)
        for k, v in _read_block(f, boff, bsize):
            entries[k.decode()] = v
    return entries


def parse_entry(v):
    pos, e = 0, {"dtype": 0, "shape": [], "offset": 0, "size": 0, "crc": None}
    while pos < len(v):
        tag, pos = _varint(v, pos)
        field, wt = tag >> 3, tag & 7
        if wt == 0:
            x, pos = _varint(v, pos)
            e[{1: "dtype", 3: "shard", 4: "offset", 5: "size"}.get(field, "x")] = x
        elif wt == 5:
            e["crc"] = struct.unpack("<I", v[pos:pos + 4])[0]
            pos += 4
        elif wt == 2:
            ln, pos = _varint(v, pos)
            sub = v[pos:pos + ln]
            pos += ln
            if field == 2:
                q = 0
                while q < len(sub):
                    t2, q = _varint(sub, q)
                    l2, q = _varint(sub, q)
                    dim = sub[q:q + l2]
                    q += l2
                    size = 0
                    if dim:
                        _, r = _varint(dim, 0)
                        size, _ = _varint(dim, r)
                    e["shape"].append(size)
    return e


def _model(seed=0, momentum=0.9):
    keras.clear_session()
    keras.layers.set_init_seed(seed)
    return build_and_compile_cnn_model(learning_rate=0.05, momentum=momentum)


def _xy(n=64, seed=0):
    rng = np.random.default_rng(seed)
    return rng.random((n, 28, 28, 1), dtype=np.float32), rng.integers(0, 10, size=n)


def test_bundle_is_readable_by_independent_reader_with_keras_names_and_layouts(tmp_path):
    m = _model()
    m.train_on_batch(*_xy())
    prefix = str(tmp_path / "ckpt" / "weights")
    m.save_weights(prefix)
    assert sorted(os.listdir(tmp_path / "ckpt")) == ["checkpoint", "weights.data-00000-of-00001", "weights.index"]
    assert latest_checkpoint(str(tmp_path / "ckpt")) == prefix
    entries = read_index(prefix + ".index")
    data = open(prefix + ".data-00000-of-00001", "rb").read()
    want = {
        "layer_with_weights-0/kernel/.ATTRIBUTES/VARIABLE_VALUE": (3, 3, 1, 32),   # HWIO, conv2d/kernel
        "layer_with_weights-0/bias/.ATTRIBUTES/VARIABLE_VALUE": (32,),
        "layer_with_weights-1/kernel/.ATTRIBUTES/VARIABLE_VALUE": (3, 3, 32, 64),
        "layer_with_weights-1/bias/.ATTRIBUTES/VARIABLE_VALUE": (64,),
        "layer_with_weights-2/kernel/.ATTRIBUTES/VARIABLE_VALUE": (1600, 128),     # (in, out), rows in (h,w,c) order
        "layer_with_weights-2/bias/.ATTRIBUTES/VARIABLE_VALUE": (128,),
        "layer_with_weights-3/kernel/.ATTRIBUTES/VARIABLE_VALUE": (128, 10),
        "layer_with_weights-3/bias/.ATTRIBUTES/VARIABLE_VALUE": (10,),
    }
    by_key = dict(keras_checkpoint_keys(m))
    assert list(entries) == sorted(entries)  # table keys are sorted
    assert "" in entries and "_CHECKPOINTABLE_OBJECT_GRAPH" in entries
    for key, shape in want.items():
        e = parse_entry(entries[key])
        assert e["dtype"] == 1 and tuple(e["shape"]) == shape  # DT_FLOAT
        raw = data[e["offset"]:e["offset"] + e["size"]]
        assert _unmask(e["crc"]) == _crc32c_py(raw)
        np.testing.assert_array_equal(np.frombuffer(raw, np.float32).reshape(shape), by_key[key].numpy())
    it = parse_entry(entries["optimizer/iter/.ATTRIBUTES/VARIABLE_VALUE"])
    assert it["dtype"] == 9 and it["shape"] == []  # DT_INT64 scalar
    assert np.frombuffer(data[it["offset"]:it["offset"] + it["size"]], np.int64)[0] == 1
    slot = "layer_with_weights-2/kernel/.OPTIMIZER_SLOT/optimizer/momentum/.ATTRIBUTES/VARIABLE_VALUE"
    assert tuple(parse_entry(entries[slot])["shape"]) == (1600, 128)
    # object graph: parse with tensorboard's own proto class
    from tensorboard.compat.proto import trackable_object_graph_pb2 as pb

    og = parse_entry(entries["_CHECKPOINTABLE_OBJECT_GRAPH"])
    blob = data[og["offset"]:og["offset"] + og["size"]]
    ln, p = _varint(blob, 0)
    g = pb.TrackableObjectGraph.FromString(blob[p + 4:p + 4 + ln])
    root_children = {c.local_name for c in g.nodes[0].children}
    assert {"layer_with_weights-0", "layer_with_weights-3", "layer-0", "layer-6", "optimizer"} <= root_children
    keys = {a.checkpoint_key for n in g.nodes for a in n.attributes}
    assert set(want) <= keys
    names = {a.full_name for n in g.nodes for a in n.attributes}
    assert {"conv2d/kernel", "conv2d_1/kernel", "dense/kernel", "dense_1/bias"} <= names


def test_roundtrip_restores_weights_and_optimizer_and_next_step_loss(tmp_path):
    x, y = _xy(128, 1)
    a = _model(1)
    a.train_on_batch(x[:64], y[:64])
    a.save_weights(str(tmp_path / "w"))
    loss_a = a.train_on_batch(x[64:], y[64:], return_dict=True)["loss"]
    b = _model(2)
    assert not np.array_equal(a.get_weights()[0], b.get_weights()[0])
    b.load_weights(str(tmp_path / "w"))
    assert b.optimizer.iterations == 1
    loss_b = b.train_on_batch(x[64:], y[64:], return_dict=True)["loss"]
    assert abs(loss_a - loss_b) < 1e-6  # resume reproduces the next step (weights AND momentum slots restored)
    for u, v in zip(a.get_weights(), b.get_weights()):
        np.testing.assert_allclose(u, v, rtol=1e-6, atol=1e-7)
    # npz second format
    a.save_weights(str(tmp_path / "w.npz"))
    c = _model(3)
    c.load_weights(str(tmp_path / "w.npz"))
    np.testing.assert_array_equal(c.get_weights()[4], a.get_weights()[4])


def test_corruption_and_shape_mismatch_are_detected(tmp_path):
    m = _model()
    prefix = str(tmp_path / "w")
    m.save_weights(prefix)
    p = prefix + ".data-00000-of-00001"
    blob = bytearray(open(p, "rb").read())
    blob[100] ^= 0xFF
    open(p, "wb").write(bytes(blob))
    with pytest.raises(Exception, match="checksum"):
        _model(1).load_weights(prefix)
    m.save_weights(prefix)
    keras.clear_session()
    other = keras.Sequential([keras.layers.Conv2D(16, 3, activation="relu", input_shape=(28, 28, 1)),
                              keras.layers.Flatten(), keras.layers.Dense(10)])
    other.compile(optimizer="sgd", loss=keras.losses.SparseCategoricalCrossentropy(from_logits=True))
    with pytest.raises(ValueError, match="shape"):
        other.load_weights(prefix)


def test_backup_and_restore_resumes_from_last_epoch(tmp_path):
    x, y = _xy(128, 4)
    m = _model(4)

    class Crash(keras.callbacks.Callback):
        def on_epoch_begin(self, epoch, logs=None):
            if epoch == 2:
                raise KeyboardInterrupt("simulated worker loss")

    bk = lambda: keras.callbacks.BackupAndRestore(str(tmp_path / "bk"))  # noqa: E731
    with pytest.raises(KeyboardInterrupt):
        m.fit(x, y, batch_size=64, epochs=4, verbose=0, shuffle=False, callbacks=[bk(), Crash()])
    w_after_2 = m.get_weights()
    m2 = _model(99)
    h = m2.fit(x, y, batch_size=64, epochs=4, verbose=0, shuffle=False, callbacks=[bk()])
    assert h.epoch == [2, 3]  # picked up after the two finished epochs
    ref = _model(4)
    ref.fit(x, y, batch_size=64, epochs=4, verbose=0, shuffle=False)
    for u, v in zip(ref.get_weights(), m2.get_weights()):
        np.testing.assert_allclose(u, v, rtol=1e-5, atol=1e-6)
    assert not os.path.exists(tmp_path / "bk" / "chief")  # cleaned up after a successful run
    assert w_after_2[0].shape == (3, 3, 1, 32)


def test_resume_restores_the_data_iterator_position(tmp_path):
    """The reference's fit(epochs, steps_per_epoch) carries ONE iterator over a shuffled dataset across epochs
    (tf_dist_example.py:31,59). A restart from a backup must continue in that iterator - same shuffle order, already
    consumed batches skipped - so the epochs after the restart see the data (and the losses) the uninterrupted run saw."""
    import tfdl

    x, y = _xy(512, 7)

    def data():
        return tfdl.data.Dataset.from_tensor_slices((x, y)).shuffle(200).batch(32)

    class Crash(keras.callbacks.Callback):
        def on_epoch_begin(self, epoch, logs=None):
            if epoch == 2:
                raise KeyboardInterrupt("simulated worker loss")

    bk = lambda: keras.callbacks.BackupAndRestore(str(tmp_path / "bk"))  # noqa: E731
    ref = _model(5)
    ds_ref = data()
    entropy = tfdl.data.dataset.shuffle_nodes(ds_ref)[0]._entropy
    h_ref = ref.fit(ds_ref, epochs=4, steps_per_epoch=3, verbose=0)
    m = _model(5)
    ds1 = data()
    tfdl.data.dataset.shuffle_nodes(ds1)[0]._entropy = entropy  # the interrupted run draws the same order as the reference
    with pytest.raises(KeyboardInterrupt):
        m.fit(ds1, epochs=4, steps_per_epoch=3, verbose=0, callbacks=[bk(), Crash()])
    m2 = _model(77)
    h2 = m2.fit(data(), epochs=4, steps_per_epoch=3, verbose=0, callbacks=[bk()])  # fresh dataset object, random seed
    assert h2.epoch == [2, 3]
    np.testing.assert_allclose(h2.history["loss"], h_ref.history["loss"][2:], rtol=1e-5)
    for u, v in zip(ref.get_weights(), m2.get_weights()):
        np.testing.assert_allclose(u, v, rtol=1e-5, atol=1e-6)


def test_tensorboard_events_are_readable_by_tensorboard(tmp_path):
    m = _model()
    x, y = _xy(64)
    m.fit(x, y, batch_size=32, epochs=2, verbose=0, callbacks=[keras.callbacks.TensorBoard(str(tmp_path / "tb"))])
    files = os.listdir(tmp_path / "tb" / "train")
    assert len(files) == 1 and files[0].startswith("events.out.tfevents.")
    from tensorboard.backend.event_processing.event_file_loader import EventFileLoader

    events = list(EventFileLoader(str(tmp_path / "tb" / "train" / files[0])).Load())
    assert events[0].file_version == "brain.Event:2"
    tags = {(v.tag, e.step) for e in events for v in e.summary.value}
    assert ("epoch_loss", 0) in tags and ("epoch_loss", 1) in tags and ("epoch_sparse_categorical_accuracy", 1) in tags


def _w_ckpt(rank, addrs, outdir):
    os.environ["TF_CONFIG"] = json.dumps({"cluster": {"worker": addrs}, "task": {"type": "worker", "index": rank}})
    import tfdl
    from tfdl import keras
    from tfdl.models import build_and_compile_cnn_model

    s = tfdl.distribute.experimental.MultiWorkerMirroredStrategy(timeout_s=30)
    with s.scope():
        m = build_and_compile_cnn_model()
    rng = np.random.default_rng(0)
    x, y = rng.random((64, 28, 28, 1), dtype=np.float32), rng.integers(0, 10, size=64)
    cb = keras.callbacks.ModelCheckpoint(os.path.join(outdir, f"rank{rank}", "ck-{epoch:02d}"))
    m.fit(x, y, batch_size=32, epochs=2, verbose=0, shuffle=False, callbacks=[cb])
    s.close()
    return sorted(os.listdir(os.path.join(outdir, f"rank{rank}"))) if os.path.isdir(os.path.join(outdir, f"rank{rank}")) else []


def test_only_the_chief_writes_checkpoints(tmp_path):
    res = run_workers(_w_ckpt, 2, args=(str(tmp_path),), timeout=120)
    assert res[0][0] == "ok" and res[1][0] == "ok", res
    assert "ck-01.index" in res[0][1] and "ck-02.data-00000-of-00001" in res[0][1] and "checkpoint" in res[0][1]
    assert res[1][1] == []  # README.md:51: the chief does the saving
