#!/usr/bin/env python
"""Host-side cost of model.fit() on the index path, without a GPU: the reference's pipeline and call
(tf_dist_example.py:20-37,59: map(scale).cache().shuffle(10000).batch(64), fit(epochs=1, steps_per_epoch=20)) driven
into a stand-in engine whose train_index_chunk / epoch_mark / epoch_read do nothing. What remains is everything
fit() itself does per call - dataset distribution, the index stream (C++ shuffle buffer + batching), callbacks, logs -
i.e. the part of a 20-step fit that no kernel can hide. Also times the C++ gather of those 20 batches, blocking and as
the background job fit() uses (time until the first two records are usable).

    python tools/fit_overhead.py [--steps 20] [--reps 300]
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

os.environ.setdefault("CUDA_VISIBLE_DEVICES", "")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402

import tfdl  # noqa: E402
from tfdl.runtime import native as rt  # noqa: E402


class NullEngine:
    """The interface fit() drives on the native path (keras/native_engine.py), with the device work left out."""
    CHUNK = 64

    def __init__(self) -> None:
        self.prof: dict = {}
        self.first_call = None
        self.groups: list = []

    def train_chunk(self, group) -> None: ...
    def train_step(self, batch) -> None: ...
    def flush_metrics(self) -> None: ...

    def index_arrays(self, base):
        return base.arrays[0], base.arrays[1]

    def train_index_chunk(self, x, y, idxs) -> None:
        if self.first_call is None:
            self.first_call = time.perf_counter()
        self.groups = idxs

    def epoch_mark(self):
        return 0

    def epoch_read(self, handle):
        return 46.0, 2.0, 1280.0


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--reps", type=int, default=300)
    a = ap.parse_args()
    rng = np.random.default_rng(0)
    x_u8 = rng.integers(0, 256, size=(60000, 28, 28, 1), dtype=np.uint8)
    y = rng.integers(0, 10, size=60000).astype(np.int64)
    ds = tfdl.data.Dataset.from_tensor_slices((x_u8, y)).map(
        lambda im, lb: (tfdl.utils.cast(im, tfdl.utils.float32) / 255, lb)).cache().shuffle(10000).batch(64)
    opts = tfdl.data.Options()
    opts.experimental_distribute.auto_shard_policy = tfdl.data.experimental.AutoShardPolicy.OFF
    ds = ds.with_options(opts)
    strategy = tfdl.distribute.MirroredStrategy(devices=["cpu"])
    with strategy.scope():
        model = tfdl.models.build_and_compile_cnn_model(learning_rate=1e-3)
    model.fit(ds, epochs=1, steps_per_epoch=2, verbose=0)  # builds the model, materialises the cache
    eng = NullEngine()
    model._engine = eng
    for _ in range(30):
        model.fit(ds, epochs=1, steps_per_epoch=a.steps, verbose=0)
    total, lead = [], []
    for _ in range(a.reps):
        eng.first_call = None
        t0 = time.perf_counter()
        model.fit(ds, epochs=1, steps_per_epoch=a.steps, verbose=0)
        total.append(time.perf_counter() - t0)
        lead.append(eng.first_call - t0)
    # the gather of the same batches
    xf, yf = eng.index_arrays(strategy.experimental_distribute_dataset(ds).index_stream()[0])
    B = 64
    rec = B * 28 * 28 * 4 + B * 4
    ring = np.zeros(a.steps * rec, dtype=np.uint8)
    sync, first2, done = [], [], []
    for _ in range(a.reps):
        flat = np.ascontiguousarray(np.concatenate(eng.groups), dtype=np.int64)
        flat = rng.permutation(60000)[:flat.size].astype(np.int64)
        t0 = time.perf_counter()
        rt.gather_records(xf, yf, flat, a.steps, B, ring.ctypes.data, rec, threads=8)
        sync.append(time.perf_counter() - t0)
        time.sleep(0.0005)
        t0 = time.perf_counter()
        job = rt.GatherJob(xf, yf, flat, a.steps, B, ring.ctypes.data, rec)
        job.wait(2)
        t1 = time.perf_counter()
        job.wait(a.steps)
        t2 = time.perf_counter()
        job.end()
        first2.append(t1 - t0)
        done.append(t2 - t0)
        time.sleep(0.0005)
    us = lambda v: round(float(np.median(v)) * 1e6, 1)  # noqa: E731
    print(json.dumps({"steps": a.steps, "cpus": os.cpu_count(),
                      "fit_python_only_us": us(total), "fit_until_first_train_call_us": us(lead),
                      "gather_blocking_us": us(sync), "gather_job_first_2_records_us": us(first2),
                      "gather_job_all_records_us": us(done)}))


if __name__ == "__main__":
    main()
