#!/usr/bin/env python
"""BASELINE.json config 1: the reference CNN on 2 localhost CPU workers (TF_CONFIG, MultiWorkerMirroredStrategy, RING).
This framework (torch CPU ops + the C++ TCP ring all-reduce of csrc/runtime/control.cc) against a stand-in
(torch CPU + DistributedDataParallel over gloo). Wall-clock, max over workers, synthetic MNIST-shaped data.

    python tools/bench_cpu_workers.py [--workers 2] [--steps 40] [--warmup 5]
"""
from __future__ import annotations

import argparse
import json
import os
import socket
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def worker_tfdl(steps: int, warmup: int) -> None:
    import numpy as np

    import tfdl

    strategy = tfdl.distribute.experimental.MultiWorkerMirroredStrategy(
        tfdl.distribute.experimental.CollectiveCommunication.RING, devices=["cpu"])
    n = strategy.num_replicas_in_sync
    rng = np.random.default_rng(0)
    G = 64 * n
    x = rng.random((G * 8, 28, 28, 1), dtype=np.float32)
    y = rng.integers(0, 10, size=G * 8).astype(np.int32)
    ds = tfdl.data.Dataset.from_tensor_slices((x, y)).batch(G).repeat()
    o = tfdl.data.Options()
    o.experimental_distribute.auto_shard_policy = tfdl.data.experimental.AutoShardPolicy.OFF
    with strategy.scope():
        model = tfdl.models.build_and_compile_cnn_model(learning_rate=1e-3)
    model.fit(ds.with_options(o), epochs=1, steps_per_epoch=warmup, verbose=0)
    strategy.cross_device_ops.barrier()
    t0 = time.perf_counter()
    model.fit(ds.with_options(o), epochs=1, steps_per_epoch=steps, verbose=0)
    strategy.cross_device_ops.barrier()
    dt = time.perf_counter() - t0
    print("CPU_BENCH " + json.dumps({"impl": "tfdl", "workers": n, "sec": dt, "images_per_sec": G * steps / dt}), flush=True)
    strategy.close()


def worker_standin(steps: int, warmup: int) -> None:
    import torch
    import torch.distributed as dist
    import torch.nn as nn
    import torch.nn.functional as F

    dist.init_process_group("gloo")
    n = dist.get_world_size()
    net = nn.Sequential(nn.Conv2d(1, 32, 3), nn.ReLU(), nn.MaxPool2d(2), nn.Conv2d(32, 64, 3), nn.ReLU(), nn.MaxPool2d(2),
                        nn.Flatten(), nn.Linear(1600, 128), nn.ReLU(), nn.Linear(128, 10))
    model = nn.parallel.DistributedDataParallel(net)
    opt = torch.optim.SGD(model.parameters(), lr=1e-3)
    g = torch.Generator().manual_seed(dist.get_rank())
    xs = torch.rand((8, 64, 1, 28, 28), generator=g)
    ys = torch.randint(0, 10, (8, 64), generator=g)

    def step(i):
        loss = F.cross_entropy(model(xs[i % 8]), ys[i % 8])
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()

    for i in range(warmup):
        step(i)
    dist.barrier()
    t0 = time.perf_counter()
    for i in range(steps):
        step(i)
    dist.barrier()
    dt = time.perf_counter() - t0
    if dist.get_rank() == 0:
        print("CPU_BENCH " + json.dumps({"impl": "standin-torch-ddp-gloo", "workers": n, "sec": dt,
                                         "images_per_sec": 64 * n * steps / dt}), flush=True)
    dist.destroy_process_group()


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--workers", type=int, default=2)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--role", default="")
    a = ap.parse_args()
    if a.role == "tfdl":
        return worker_tfdl(a.steps, a.warmup)
    if a.role == "standin":
        return worker_standin(a.steps, a.warmup)
    threads = max(1, (os.cpu_count() or 2) // a.workers)
    base_env = {k: v for k, v in os.environ.items() if k not in ("TF_CONFIG", "RANK", "WORLD_SIZE", "LOCAL_RANK")}
    base_env.update(OMP_NUM_THREADS=str(threads), MKL_NUM_THREADS=str(threads), CUDA_VISIBLE_DEVICES="")
    out = []
    addrs = [f"127.0.0.1:{free_port()}" for _ in range(a.workers)]
    procs = []
    for i in range(a.workers):
        env = dict(base_env, TF_CONFIG=json.dumps({"cluster": {"worker": addrs}, "task": {"type": "worker", "index": i}}))
        procs.append(subprocess.Popen([sys.executable, __file__, "--role", "tfdl", "--steps", str(a.steps), "--warmup",
                                       str(a.warmup)], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    res = [p.communicate(timeout=1200)[0] for p in procs]
    out += [json.loads(l[len("CPU_BENCH "):]) for r in res for l in r.splitlines() if l.startswith("CPU_BENCH ")]
    port = free_port()
    procs = []
    for i in range(a.workers):
        env = dict(base_env, RANK=str(i), WORLD_SIZE=str(a.workers), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, __file__, "--role", "standin", "--steps", str(a.steps), "--warmup",
                                       str(a.warmup)], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    res = [p.communicate(timeout=1200)[0] for p in procs]
    out += [json.loads(l[len("CPU_BENCH "):]) for r in res for l in r.splitlines() if l.startswith("CPU_BENCH ")]
    ours = min((o["images_per_sec"] for o in out if o["impl"] == "tfdl"), default=None)
    ref = next((o["images_per_sec"] for o in out if o["impl"] != "tfdl"), None)
    print(json.dumps({"config": "mnist_cnn, 2 localhost CPU workers, global batch 128", "threads_per_worker": threads,
                      "tfdl_images_per_sec": ours, "standin_images_per_sec": ref, "raw": out}))
    if ours is None or ref is None:
        print("\n".join(res)[-3000:])


if __name__ == "__main__":
    main()
