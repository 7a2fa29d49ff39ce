#!/usr/bin/env python
"""The headline step (reference CNN, 64 images per GPU) with MirroredStrategy() in ONE plain python process driving
every visible GPU - the launch form /root/reference/tf_dist_example.py:13 names - for comparison with the
one-process-per-GPU launch that bench.py times. Device-timed with CUDA events on every GPU's stream, max over GPUs.

    python tools/bench_single_process.py [--gpus N] [--steps K] [--warmup W]"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
    os.environ.pop(k, None)
import tfdl  # noqa: E402
from tfdl.models import build_and_compile_cnn_model  # noqa: E402


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=0)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=16)
    a = ap.parse_args()
    n = a.gpus or torch.cuda.device_count()
    strategy = tfdl.distribute.MirroredStrategy(devices=[f"gpu:{i}" for i in range(n)])
    with strategy.scope():
        model = build_and_compile_cnn_model(learning_rate=1e-3)
    eng = model._make_engine()
    eng.prepare(64)
    reps = eng.replicas if hasattr(eng, "replicas") else [eng]
    rec = eng.rec_bytes
    n_pool = max(16, (160 << 20) // rec)  # larger than L2 on every GPU
    pools = []
    for r, rep in enumerate(reps):
        rng = np.random.default_rng(1234 + r)
        hp = np.empty((n_pool, rec), np.uint8)
        nx = 64 * 28 * 28 * 4
        hp[:, :nx].view(np.float32)[:] = rng.random((n_pool, nx // 4), dtype=np.float32)
        hp[:, nx:].view(np.int32)[:] = rng.integers(0, 10, size=(n_pool, 64), dtype=np.int32)
        pools.append(torch.from_numpy(hp).to(rep.device))
    run = (lambda k: eng.run_device(pools, k)) if hasattr(eng, "replicas") else (lambda k: eng.run_device(pools[0], k))
    run(max(a.warmup, 8))
    for rep in reps:
        rep.stream.synchronize()
    ev = []
    for rep in reps:
        with torch.cuda.device(rep.device):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(rep.stream)
            ev.append((e0, e1))
    run(a.steps)
    for rep, (_, e1) in zip(reps, ev):
        with torch.cuda.device(rep.device):
            e1.record(rep.stream)
    for rep in reps:
        rep.stream.synchronize()
    ms = max(e0.elapsed_time(e1) for e0, e1 in ev)
    print("SINGLE_PROCESS " + json.dumps({"n_gpus": n, "engine": type(eng).__name__, "allreduce": eng.algo_name,
                                          "steps": a.steps, "ms_per_step": ms / a.steps,
                                          "images_per_sec": 64 * n * a.steps / (ms / 1e3)}), flush=True)


if __name__ == "__main__":
    main()
