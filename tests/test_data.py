"""Dataset API: the calls the reference makes plus shard policies and the global->replica split."""
import numpy as np
import pytest

import tfdl
from tfdl.data import AutoShardPolicy, Dataset, Options, auto_shard, rebatch_for_replicas, datasets
from tfdl.data.dataset import _shuffle_indices_numpy


def _xy(n=100):
    return np.arange(n * 4, dtype=np.uint8).reshape(n, 2, 2, 1), np.arange(n, dtype=np.int64)


def test_reference_pipeline_shapes_and_scaling():
    ds, info = datasets.load(name="mnist", with_info=True, as_supervised=True, synthetic_size=(1000, 100))
    assert info.splits["train"].num_examples == 1000 and info.features["image"][0] == (28, 28, 1)

    def scale(image, label):
        image = tfdl.utils.cast(image, tfdl.utils.float32)
        image /= 255
        return image, label

    train = ds["train"].map(scale).cache().shuffle(10000).batch(128)
    opt = Options()
    opt.experimental_distribute.auto_shard_policy = AutoShardPolicy.OFF
    train = train.with_options(opt)
    batches = list(train)
    assert len(batches) == 8 and train.cardinality() == 8
    x, y = batches[0]
    assert x.shape == (128, 28, 28, 1) and x.dtype == np.float32 and 0 <= x.min() and x.max() <= 1
    assert y.shape == (128,) and batches[-1][0].shape[0] == 1000 - 7 * 128
    assert train.options().experimental_distribute.auto_shard_policy == AutoShardPolicy.OFF
    seen = np.sort(np.concatenate([b[1] for b in batches]))
    ref = np.sort(np.concatenate([b[1] for b in ds["train"].batch(1000)]))
    np.testing.assert_array_equal(seen, ref)  # one epoch = every element exactly once


def test_map_vectorized_probe_and_per_element_fallback():
    x, y = _xy()
    calls = []

    def per_elem(a, b):
        calls.append(1)
        assert a.shape == (2, 2, 1)  # raises on a batch -> per-element path
        return a.astype(np.float32).sum(), b

    out = list(Dataset.from_tensor_slices((x, y)).map(per_elem).batch(10))
    assert out[0][0].shape == (10,) and len(calls) > 10
    np.testing.assert_allclose(out[0][0], x[:10].astype(np.float32).reshape(10, -1).sum(1))
    vec = Dataset.from_tensor_slices((x, y)).map(lambda a, b: (a.astype(np.float32) / 255, b))
    assert vec._base().vectorized


def test_cache_freezes_parent_and_runs_map_once():
    x, y = _xy(20)
    n = [0]

    def f(a, b):
        n[0] += 1
        return a * 2, b

    ds = Dataset.from_tensor_slices((x, y)).map(f, vectorized=True).cache()
    list(ds.batch(5)), list(ds.batch(5))
    first = n[0]
    list(ds.batch(5))
    assert n[0] == first  # served from the cache


def test_shuffle_semantics_and_reshuffle_each_iteration():
    src = np.arange(1000)
    out = _shuffle_indices_numpy(src, 50, np.random.default_rng(0))
    assert sorted(out.tolist()) == src.tolist()
    assert np.all(out[:950] < np.arange(950) + 51)  # buffer locality
    ds = Dataset.range(200).shuffle(200, seed=3)
    a, b = np.array(list(ds)), np.array(list(ds))
    assert sorted(a.tolist()) == list(range(200)) and not np.array_equal(a, b)
    fixed = Dataset.range(200).shuffle(200, seed=3, reshuffle_each_iteration=False)
    assert np.array_equal(np.array(list(fixed)), np.array(list(fixed)))


def test_batch_repeat_take_skip_shard_unbatch():
    ds = Dataset.range(10)
    assert [b.tolist() for b in ds.batch(4)] == [[0, 1, 2, 3], [4, 5, 6, 7], [8, 9]]
    assert [b.tolist() for b in ds.batch(4, drop_remainder=True)] == [[0, 1, 2, 3], [4, 5, 6, 7]]
    assert len(list(ds.repeat(3))) == 30 and ds.repeat().cardinality() == -1
    assert list(ds.take(3)) == [0, 1, 2] and list(ds.skip(8)) == [8, 9]
    assert list(ds.shard(3, 1)) == [1, 4, 7] and ds.shard(3, 1).cardinality() == 3
    assert list(ds.batch(4).unbatch()) == list(range(10))
    assert [b.tolist() for b in ds.batch(4).rebatch(5)] == [[0, 1, 2, 3, 4], [5, 6, 7, 8, 9]]
    assert list(ds.batch(2).prefetch(2))[0].tolist() == [0, 1]
    with pytest.raises(ValueError):
        ds.shard(2, 2)


def test_generic_path_matches_indexable_path():
    gen = Dataset.from_generator(lambda: iter(range(10)))
    assert [b.tolist() for b in gen.batch(4)] == [[0, 1, 2, 3], [4, 5, 6, 7], [8, 9]]
    assert sorted(np.concatenate(list(gen.shuffle(4, seed=1).batch(3))).tolist()) == list(range(10))
    assert list(gen.shard(2, 0)) == [0, 2, 4, 6, 8] and list(gen.repeat(2).take(12))[-1] == 1


def test_rebatch_global_to_per_replica():
    x, y = _xy(64)
    ds = Dataset.from_tensor_slices((x, y)).batch(16)
    re, per = rebatch_for_replicas(ds, 2)  # GLOBAL_BATCH_SIZE = per-replica * replicas (tf_dist_example.py:18)
    assert per == 8 and [b[1].tolist() for b in re][:2] == [list(range(8)), list(range(8, 16))]
    with pytest.raises(ValueError):
        rebatch_for_replicas(ds, 3)
    with pytest.raises(ValueError):
        rebatch_for_replicas(Dataset.range(4), 2)


def test_auto_shard_policies():
    x, y = _xy(40)
    base = Dataset.from_tensor_slices((x, y)).batch(8)

    def labels(d):
        return np.concatenate([b[1] for b in d]).tolist()

    off = Options()
    off.experimental_distribute.auto_shard_policy = AutoShardPolicy.OFF
    assert labels(auto_shard(base.with_options(off), 2, 1)) == list(range(40))  # every worker sees everything
    data = Options()
    data.experimental_distribute.auto_shard_policy = AutoShardPolicy.DATA
    s0, s1 = labels(auto_shard(base.with_options(data), 2, 0)), labels(auto_shard(base.with_options(data), 2, 1))
    assert s0 == list(range(0, 40, 2)) and s1 == list(range(1, 40, 2))
    assert labels(auto_shard(base, 2, 0)) == s0  # AUTO falls back to DATA for in-memory sources
    f = Options()
    f.experimental_distribute.auto_shard_policy = AutoShardPolicy.FILE
    with pytest.raises(ValueError, match="FILE"):
        auto_shard(base.with_options(f), 2, 0)
    # single-file source with 1 worker is fine
    assert labels(auto_shard(base.with_options(f), 1, 0)) == list(range(40))


def test_idx_reader_roundtrip(tmp_path):
    import gzip
    import struct

    d = tmp_path / "mnist"
    d.mkdir()
    rng = np.random.default_rng(0)
    for stem, n in (("train", 30), ("t10k", 10)):
        img = rng.integers(0, 255, (n, 28, 28), dtype=np.uint8)
        lab = rng.integers(0, 10, n, dtype=np.uint8)
        with gzip.open(d / f"{stem}-images-idx3-ubyte.gz", "wb") as f:
            f.write(struct.pack(">IIII", 0x803, n, 28, 28) + img.tobytes())
        with open(d / f"{stem}-labels-idx1-ubyte", "wb") as f:
            f.write(struct.pack(">II", 0x801, n) + lab.tobytes())
    ds, info = datasets.load("mnist", with_info=True, as_supervised=True, data_dir=str(tmp_path))
    assert not info.synthetic and info.splits["train"].num_examples == 30
    x, y = next(iter(ds["train"].batch(30)))
    assert x.shape == (30, 28, 28, 1) and x.dtype == np.uint8 and y.dtype == np.int64


def test_index_stream_matches_the_materialised_batches():
    """The native executor gathers its batches itself from the dataset arrays; the row numbers it is handed must describe
    exactly the batches that iterating the dataset yields (tf_dist_example.py:31-33: map -> cache -> shuffle -> batch),
    also when a step is split over several local replicas."""
    import tfdl
    from tfdl.data import dataset as D

    x = np.arange(200 * 4, dtype=np.float32).reshape(200, 4)
    y = np.arange(200, dtype=np.int64)
    ds = tfdl.data.Dataset.from_tensor_slices((x, y)).map(lambda a, b: (a * 2, b)).cache().shuffle(50, seed=7).batch(16)
    base, it = D.index_stream(ds)
    idx_batches = list(it)
    xs, ys = base.arrays
    ds2 = tfdl.data.Dataset.from_tensor_slices((x, y)).map(lambda a, b: (a * 2, b)).cache().shuffle(50, seed=7).batch(16)
    real = list(ds2)
    assert len(idx_batches) == len(real) == 13
    for idx, (bx, by) in zip(idx_batches, real):
        np.testing.assert_array_equal(xs[idx], bx)
        np.testing.assert_array_equal(ys[idx], by)
    # two local replicas: one step = two consecutive per-replica batches, concatenated in replica order
    s = tfdl.distribute.MirroredStrategy(devices=["cpu:0", "cpu:1"])
    ds3 = tfdl.data.Dataset.from_tensor_slices((x, y)).batch(16)
    dist = s.experimental_distribute_dataset(ds3)
    base3, steps = dist.index_stream()
    steps = list(steps)
    assert [len(i) for i in steps] == [16] * 12 + [8]
    np.testing.assert_array_equal(steps[0], np.arange(16))
    for st, per in zip(steps, dist):
        np.testing.assert_array_equal(np.concatenate([np.asarray(b[1]) for b in per]), y[st])


def test_index_stream_equals_iteration_for_random_pipelines():
    """Property test (hypothesis): for pipelines assembled at random from the indexable stages - shuffle (seeded), repeat,
    take, skip, shard, map, cache - the row numbers of `index_stream` gather exactly the batches that iterating the same
    pipeline yields, including the ragged last batch and drop_remainder, and also after auto-sharding for a worker."""
    from hypothesis import HealthCheck, given, settings
    from hypothesis import strategies as st

    import tfdl
    from tfdl.data import dataset as D

    stage = st.one_of(
        st.tuples(st.just("shuffle"), st.integers(1, 64), st.integers(0, 1000)),
        st.tuples(st.just("repeat"), st.integers(1, 3)),
        st.tuples(st.just("take"), st.integers(0, 90)),
        st.tuples(st.just("skip"), st.integers(0, 40)),
        st.tuples(st.just("shard"), st.integers(1, 4), st.integers(0, 3)),
        st.tuples(st.just("map")),
        st.tuples(st.just("cache")),
    )

    def build(n, stages, bs, drop):
        x = (np.arange(n * 3, dtype=np.float32).reshape(n, 3) + 0.5)
        y = np.arange(n, dtype=np.int64)
        ds = tfdl.data.Dataset.from_tensor_slices((x, y))
        for s in stages:
            if s[0] == "shuffle":
                ds = ds.shuffle(s[1], seed=s[2], reshuffle_each_iteration=False)
            elif s[0] == "repeat":
                ds = ds.repeat(s[1])
            elif s[0] == "take":
                ds = ds.take(s[1])
            elif s[0] == "skip":
                ds = ds.skip(s[1])
            elif s[0] == "shard":
                ds = ds.shard(s[1], s[2] % s[1])
            elif s[0] == "map":
                ds = ds.map(lambda a, b: (a * 2 + 1, b))
            elif s[0] == "cache":
                ds = ds.cache()
        return ds.batch(bs, drop_remainder=drop)

    @settings(max_examples=60, deadline=None, suppress_health_check=[HealthCheck.too_slow])
    @given(n=st.integers(1, 80), stages=st.lists(stage, max_size=5), bs=st.integers(1, 17), drop=st.booleans(),
           workers=st.integers(1, 3))
    def check(n, stages, bs, drop, workers):
        ds = build(n, stages, bs, drop)
        real = list(build(n, stages, bs, drop))
        st_ = D.index_stream(ds)
        if st_ is None:
            return  # not an indexable pipeline: fit() iterates it instead
        base, it = st_
        got = list(it)
        arrays = base.arrays if hasattr(base, "arrays") else None
        assert len(got) == len(real)
        for idx, (bx, by) in zip(got, real):
            e = base.gather(idx)
            np.testing.assert_array_equal(e[0], bx)
            np.testing.assert_array_equal(e[1], by)
            if arrays is not None:
                np.testing.assert_array_equal(arrays[1][idx], by)
        if workers > 1:  # the DATA policy on top: worker w's stream is what worker w's sharded dataset yields
            for w in range(workers):
                sharded = D.auto_shard(build(n, stages, bs, drop), workers, w, tfdl.data.AutoShardPolicy.DATA)
                want = list(D.auto_shard(build(n, stages, bs, drop), workers, w, tfdl.data.AutoShardPolicy.DATA))
                sw = D.index_stream(sharded)
                if sw is None:
                    continue
                bw, iw = sw
                gotw = list(iw)
                assert len(gotw) == len(want)
                for idx, (bx, by) in zip(gotw, want):
                    np.testing.assert_array_equal(bw.gather(idx)[1], by)

    check()
