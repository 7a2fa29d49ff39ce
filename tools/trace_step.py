#!/usr/bin/env python
"""Where a training step of the reference CNN spends its time, per rank (marker kernels between the stages,
csrc/kernels/program.cu:trace_mark). Launch like bench.py (torchrun for N > 1):

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/trace_step.py [--steps 400]

Prints, for every rank, the mean time of each marker relative to the step's start marker and the mean step period;
rank 0 writes gpurun_out/trace_step_N<N>.json."""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tfdl  # noqa: E402
from tfdl.models import build_and_compile_cnn_model  # noqa: E402
from tfdl.ops import _lib  # noqa: E402

NAMES = ["start", "conv_fwd", "dense_block", "conv2_wgrad(side)", "conv_bwd", "dense_AR(side)", "conv_AR=end"]
CAP = 256


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=600)
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    strategy = tfdl.distribute.MirroredStrategy() if world > 1 else tfdl.distribute.MirroredStrategy(devices=["gpu:0"])
    dev = strategy.local_devices[0]
    with strategy.scope():
        model = build_and_compile_cnn_model(learning_rate=1e-3)
    buf = torch.zeros(8 * CAP, dtype=torch.int64, device=dev)
    cnt = torch.zeros(8, dtype=torch.int32, device=dev)
    lib = _lib.kernels()
    lib.tfdl_trace_enable.argtypes = [C.c_void_p, C.c_void_p]
    lib.tfdl_trace_enable(buf.data_ptr(), cnt.data_ptr())  # before the step graphs are captured
    engine = model._make_engine()
    engine.prepare(64)
    rec = engine.rec_bytes
    pool = torch.randint(0, 255, (1024, rec), dtype=torch.uint8)
    nx = 64 * 28 * 28 * 4
    pool[:, :nx].view(torch.float32)[:] = torch.rand(1024, nx // 4)
    pool[:, nx:].view(torch.int32)[:] = torch.randint(0, 10, (1024, 64), dtype=torch.int32)
    dpool = pool.to(dev)
    import time
    engine.run_device(dpool, args.steps)
    engine.stream.synchronize()
    t0 = time.perf_counter()
    engine.run_device(dpool, 24)          # few enough to fit the launch queue: how long does the HOST need per step?
    host_us = (time.perf_counter() - t0) / 24 * 1e6
    engine.stream.synchronize()
    t = buf.cpu().numpy().reshape(8, CAP).astype(np.float64)
    n = int(cnt.cpu()[0])
    k = min(n, CAP) - 8
    order = np.argsort(t[0, :min(n, CAP)])[-k:]  # the last k steps, in time order
    cn = cnt.cpu().numpy()
    rel = {NAMES[i]: float(np.mean(t[i, order] - t[0, order]) / 1e3) for i in range(1, 7) if cn[i] > 0}
    starts = np.sort(t[0, order])
    period = float(np.mean(np.diff(starts)) / 1e3)
    out = {"rank": rank, "world": world, "algo": engine.algo_name, "us_after_step_start": rel, "step_period_us": period,
           "host_enqueue_us_per_step": host_us}
    if world > 1:
        import torch.distributed as dist

        allo = [None] * world
        dist.all_gather_object(allo, out)
    else:
        allo = [out]
    if rank == 0:
        for o in allo:
            print("TRACE " + json.dumps(o), flush=True)
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        json.dump(allo, open(os.path.join(ROOT, "gpurun_out", f"trace_step_N{world}.json"), "w"), indent=1)
    strategy.close()


if __name__ == "__main__":
    main()
