#!/usr/bin/env python
"""Headline benchmark: images/sec of the reference CNN training step (BASELINE.json config 2).

    python bench.py --gpus N --steps K --warmup W          # N = 1; N > 1 is launched by torchrun
    python bench.py --impl reference ...                     # the unmodified reference (TensorFlow): unavailable here
    python bench.py --impl standin ...                       # torch + cuDNN/cuBLAS + NCCL in one CUDA graph, same model/config (labelled stand-in)

Model/config: MNIST-shape CNN of /root/reference/tf_dist_example.py:40-52 (Conv32-Pool-Conv64-Pool-
Dense128-Dense10, SGD lr=1e-3, sparse softmax cross-entropy), MirroredStrategy, global batch 64 x N
(weak scaling, 64 per GPU), synthetic 28x28x1 data, random-init weights, bf16 tensor-core operands with
fp32 accumulation and fp32 master weights.

Two numbers:
  value      device-timed (CUDA events, max over ranks): K full training steps (forward, backward,
             fused all-reduce + SGD), inputs rotating through a device-resident pool larger than L2.
  e2e.value  the same metric through the public API - model.fit(dataset, steps_per_epoch=K) - with
             the per-step host->device copy of the batch from pinned memory and the per-step
             device->host copy of (loss, accuracy) inside the timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "images/sec (device-timed, max over ranks)"
MODEL = "mnist_cnn_conv32_conv64_dense128 (tf_dist_example.py:40-52), SGD lr=1e-3"


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons every 200 ms while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int) -> None:
        self.gpu, self.rows, self.proc = gpu_index, [], None

    def start(self) -> None:
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self) -> None:
        assert self.proc is not None and self.proc.stdout is not None
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def _env_world():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def _max_over_ranks(ms: float, world: int, dev) -> float:
    if world == 1:
        return ms
    import torch
    import torch.distributed as dist

    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def _barrier(world: int, dev) -> None:
    import torch

    if world > 1:
        import torch.distributed as dist

        dist.barrier()
    torch.cuda.synchronize(dev)


def run_ours(args) -> dict:
    import numpy as np
    import torch

    import tfdl
    from tfdl.models import build_and_compile_cnn_model

    rank, world, local_rank = _env_world()
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch N>1 with torch.distributed.run")
    B = args.batch_per_gpu
    strategy = tfdl.distribute.MirroredStrategy() if world > 1 else tfdl.distribute.MirroredStrategy(devices=["gpu:0"])
    dev = strategy.local_devices[0]
    with strategy.scope():
        model = build_and_compile_cnn_model(learning_rate=1e-3)
    engine = model._make_engine()
    assert type(engine).__name__ == "NativeCnnEngine", f"native engine not selected: {type(engine).__name__}"
    engine.prepare(B)

    # ---- synthetic data: a pool of distinct batches larger than L2 (126 MB) ------------------------------
    rec = engine.rec_bytes
    n_pool = max(16, (160 << 20) // rec)
    rng = np.random.default_rng(1234 + rank)
    host_pool = torch.empty((n_pool, rec), dtype=torch.uint8).pin_memory()
    hp = host_pool.numpy()
    nx = B * 28 * 28 * 4
    hp[:, :nx].view(np.float32)[:] = rng.random((n_pool, B * 28 * 28), dtype=np.float32)
    hp[:, nx:].view(np.int32)[:] = rng.integers(0, 10, size=(n_pool, B), dtype=np.int32)
    dev_pool = host_pool.to(dev)

    # ---- kernel-only, device timed --------------------------------------------------------------------
    stream = engine.stream
    engine.run_device(dev_pool, max(args.warmup, 3))
    _barrier(world, dev)
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    _barrier(world, dev)
    # 16 more untimed steps queued directly in front of the timed ones: the ranks' all-reduces make them run in lockstep on
    # the device, so the timed K steps do not include the host-side launch skew that follows a barrier (a K = 20 run is only
    # 0.8 ms long). Still exactly K steps between the two events.
    engine.run_device(dev_pool, 16)
    e0.record(stream)
    engine.run_device(dev_pool, args.steps)
    e1.record(stream)
    _barrier(world, dev)
    ms = _max_over_ranks(e0.elapsed_time(e1), world, dev)
    clocks = sampler.stop()
    engine.flush_metrics()

    # ---- end to end through the public API: the reference's own pipeline and fit() call -------------------------
    # tf_dist_example.py:20-37,59: MNIST-sized uint8 images -> map(cast/255) -> cache -> shuffle(10000) -> batch(GLOBAL),
    # AutoShardPolicy.OFF, fit(epochs=E, steps_per_epoch=20). K timed steps = K/20 epochs of 20 steps when 20 divides K.
    n_img = 60000
    rng2 = np.random.default_rng(99)  # same data on every worker (policy OFF: every worker sees the whole set)
    x_u8 = rng2.integers(0, 256, size=(n_img, 28, 28, 1), dtype=np.uint8)
    y_i64 = rng2.integers(0, 10, size=(n_img,), dtype=np.int64)

    def scale(image, label):
        return tfdl.utils.cast(image, tfdl.utils.float32) / 255, label

    ds = tfdl.data.Dataset.from_tensor_slices((x_u8, y_i64)).map(scale).cache().shuffle(10000)
    if args.steps * B > n_img:
        ds = ds.repeat()  # the reference never runs out (200 x 64 < 60000); longer runs need another pass
    ds = ds.batch(B * world)
    opts = tfdl.data.Options()
    opts.experimental_distribute.auto_shard_policy = tfdl.data.AutoShardPolicy.OFF
    ds = ds.with_options(opts)
    spe = 20 if args.steps % 20 == 0 else args.steps
    model.fit(ds, epochs=1, steps_per_epoch=max(args.warmup, 3), verbose=0)  # also materialises the cache
    _barrier(world, dev)
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    f0.record(stream)
    hist = model.fit(ds, epochs=args.steps // spe, steps_per_epoch=spe, verbose=0)
    f1.record(stream)
    _barrier(world, dev)
    wall_ms = (time.perf_counter() - t0) * 1e3
    e2e_ms = _max_over_ranks(max(f0.elapsed_time(f1), wall_ms), world, dev)

    global_batch = B * world
    value = global_batch * args.steps / (ms / 1e3)
    out = {
        "metric": METRIC, "value": round(value, 1), "unit": "images/sec", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": round(ms / args.steps, 5), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"model": MODEL, "global_batch": global_batch, "per_gpu_batch": B, "image": "28x28x1",
                   "parallelism": f"dp{world}", "allreduce": engine.algo_name,
                   "l2_policy": f"inputs larger than L2: {n_pool} distinct device-resident batches "
                                f"({n_pool * rec >> 20} MiB) rotated, one D2D copy per step on the copy stream (overlaps the previous step, like the H2D copy of the e2e path)",
                   "optimizer_step_in_timed_region": True},
        "clocks": clocks,
        "e2e": {"value": round(global_batch * args.steps / (e2e_ms / 1e3), 1), "unit": "images/sec",
                "ms_per_step": round(e2e_ms / args.steps, 5),
                "api": f"model.fit(dataset, epochs={args.steps // spe}, steps_per_epoch={spe})",
                "pipeline": "from_tensor_slices(uint8 60000x28x28x1).map(cast/255).cache().shuffle(10000).batch(GLOBAL), "
                            "AutoShardPolicy.OFF (tf_dist_example.py:20-37)",
                "timed": "wall clock around fit() (>= the CUDA-event time), max over ranks",
                "h2d_bytes_per_step": rec, "d2h_bytes_per_step": 8,
                "final_loss": float(hist.history["loss"][-1])},
        "gpu_launches": int(args.steps * engine.launches_per_step),
        "impl": "tfdl",
    }
    strategy.close()
    return out if rank == 0 else {}


MODEL_CFG = {
    "resnet50": {"metric": "images/sec (device-timed, max over ranks)", "unit": "images/sec", "batch": 256,
                 "name": "Keras ResNet-50 (25,636,712 parameters), SGD momentum 0.9, synthetic 224x224x3"},
    "bert": {"metric": "sequences/sec (device-timed, max over ranks)", "unit": "sequences/sec", "batch": 64,
             "name": "BERT-base MLM (12 layers, hidden 768, vocab 30522), SGD, synthetic sequences of 128 tokens"},
}


def run_model(args) -> dict:
    """BASELINE.json configs 3 and 4 through the same contract: `--model resnet50|bert`, bf16 (mixed_bfloat16 policy),
    device-timed steps on device-resident synthetic batches, and model.fit() on host arrays for the end-to-end number."""
    import numpy as np
    import torch

    import tfdl
    from tfdl import keras
    from tfdl.parallel.values import PerReplica

    cfg = MODEL_CFG[args.model]
    rank, world, local_rank = _env_world()
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch N>1 with torch.distributed.run")
    B = args.batch_per_gpu or cfg["batch"]
    keras.mixed_precision.set_global_policy("mixed_bfloat16")
    strategy = tfdl.distribute.MirroredStrategy() if world > 1 else tfdl.distribute.MirroredStrategy(devices=["gpu:0"])
    dev = strategy.local_devices[0]
    with strategy.scope():
        if args.model == "resnet50":
            from tfdl.models import build_and_compile_resnet50

            model = build_and_compile_resnet50(0.1, 0.9)
        else:
            from tfdl.models import build_and_compile_bert_mlm

            model = build_and_compile_bert_mlm(1e-3, optimizer="sgd")
    rng = np.random.default_rng(7 + rank)
    n_host = 4  # distinct host batches (ResNet: 154 MB each, so every step's input comes from HBM, not L2)
    if args.model == "resnet50":
        xs = [rng.random((B, 224, 224, 3), dtype=np.float32) for _ in range(n_host)]
        ys = [rng.integers(0, 1000, size=(B,)) for _ in range(n_host)]
    else:
        xs = [rng.integers(0, 30522, size=(B, 128)) for _ in range(n_host)]
        ys = [np.where(rng.random((B, 128)) < 0.15, x, -100) for x in xs]
    eng = model._make_engine()
    dbatches = [PerReplica([(torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev))]) for x, y in zip(xs, ys)]
    W = max(args.warmup, 5)  # three eager steps precede the step graph's capture
    for i in range(W):
        eng.train_step(dbatches[i % n_host])
    _barrier(world, dev)
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    _barrier(world, dev)
    e0.record()
    for i in range(args.steps):
        eng.train_step(dbatches[i % n_host])
    e1.record()
    _barrier(world, dev)
    ms = _max_over_ranks(e0.elapsed_time(e1), world, dev)
    clocks = sampler.stop()
    # ---- end to end: fit() on host arrays, one pinned H2D copy of the batch per step, loss read back at the end ----
    xh, yh = np.concatenate(xs), np.concatenate(ys)
    reps = -(-args.steps // n_host)
    ds = tfdl.data.Dataset.from_tensor_slices((xh, yh)).batch(B * world).repeat(reps + 1)
    opts = tfdl.data.Options()
    opts.experimental_distribute.auto_shard_policy = tfdl.data.AutoShardPolicy.OFF
    ds = ds.with_options(opts)
    model.fit(ds, epochs=1, steps_per_epoch=2, verbose=0)
    _barrier(world, dev)
    t0 = time.perf_counter()
    hist = model.fit(ds, epochs=1, steps_per_epoch=args.steps, verbose=0)
    _barrier(world, dev)
    e2e_ms = _max_over_ranks((time.perf_counter() - t0) * 1e3, world, dev)
    gb = B * world
    fused = getattr(eng, "_fused", None)
    out = {
        "metric": cfg["metric"], "value": round(gb * args.steps / (ms / 1e3), 1), "unit": cfg["unit"], "n_gpus": world,
        "steps": args.steps, "warmup": W, "ms_per_step": round(ms / args.steps, 4), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"model": cfg["name"], "global_batch": gb, "per_gpu_batch": B, "parallelism": f"dp{world}",
                   "engine": type(eng).__name__, "fused_allreduce_sgd": fused is not None,
                   "buckets": getattr(eng, "bucket_info", None),
                   "l2_policy": "inputs and activations far larger than L2 (126 MB): every step streams from HBM",
                   "optimizer_step_in_timed_region": True},
        "clocks": clocks,
        "e2e": {"value": round(gb * args.steps / (e2e_ms / 1e3), 1), "unit": cfg["unit"],
                "ms_per_step": round(e2e_ms / args.steps, 4), "api": "model.fit(dataset, steps_per_epoch=K)",
                "h2d_bytes_per_step": int(xs[0].nbytes + ys[0].nbytes), "d2h_bytes_per_step": 8,
                "final_loss": float(hist.history["loss"][-1])},
        "gpu_launches": int(getattr(eng, "launches_per_step", 0) * args.steps),
        "impl": "tfdl",
    }
    strategy.close()
    return out if rank == 0 else {}


def run_standin(args) -> dict:
    """Stand-in baseline (NOT the reference): the same model on PyTorch library kernels (cuDNN / cuBLAS, bf16 autocast,
    channels_last) with an NCCL all-reduce of one flat gradient bucket, the WHOLE step (zero grads, forward, backward,
    all-reduce, SGD) captured in one CUDA graph and replayed - the competent way to run a 1-GFLOP step on torch, and the
    analogue of TensorFlow running it from a traced tf.function. Inputs come from pinned host memory every step."""
    import numpy as np
    import torch
    import torch.nn as nn
    import torch.nn.functional as F

    rank, world, local_rank = _env_world()
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=dev)
    B = args.batch_per_gpu
    net = nn.Sequential(nn.Conv2d(1, 32, 3), nn.ReLU(), nn.MaxPool2d(2), nn.Conv2d(32, 64, 3), nn.ReLU(), nn.MaxPool2d(2),
                        nn.Flatten(), nn.Linear(1600, 128), nn.ReLU(), nn.Linear(128, 10)).to(dev)
    net = net.to(memory_format=torch.channels_last)
    params = list(net.parameters())
    n_tot = sum(p.numel() for p in params)
    flat_p, flat_g = torch.zeros(n_tot, device=dev), torch.zeros(n_tot, device=dev)
    off = 0
    for p in params:  # parameters and gradients become views of two flat buffers: one all-reduce, one update
        n = p.numel()
        flat_p[off:off + n].copy_(p.data.reshape(-1))
        p.data = flat_p[off:off + n].view(p.shape)
        p.grad = flat_g[off:off + n].view(p.shape)
        off += n
    if world > 1:
        dist.broadcast(flat_p, src=0)
    lr = 1e-3
    n_pool = 256
    g = torch.Generator(device="cpu").manual_seed(rank)
    hx = torch.rand((n_pool, B, 1, 28, 28), generator=g).pin_memory()
    hy = torch.randint(0, 10, (n_pool, B), generator=g).pin_memory()
    sx = torch.zeros((B, 1, 28, 28), device=dev).contiguous(memory_format=torch.channels_last)
    sy = torch.zeros((B,), dtype=torch.long, device=dev)
    sloss = torch.zeros((), device=dev)

    def body():
        flat_g.zero_()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            loss = F.cross_entropy(net(sx), sy)
        loss.backward()
        if world > 1:
            dist.all_reduce(flat_g)
        flat_p.add_(flat_g, alpha=-lr / world)
        sloss.copy_(loss.detach())

    side = torch.cuda.Stream(dev)
    with torch.cuda.stream(side):
        for _ in range(3):
            body()
    side.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=side):
        body()
    cur = torch.cuda.current_stream(dev)
    out_host = torch.zeros(max(args.steps, args.warmup, 3)).pin_memory()

    def step(i):
        sx.copy_(hx[i % n_pool], non_blocking=True)
        sy.copy_(hy[i % n_pool], non_blocking=True)
        graph.replay()
        out_host[i:i + 1].copy_(sloss.reshape(1), non_blocking=True)

    for i in range(max(args.warmup, 3)):
        step(i)
    _barrier(world, dev)
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(cur)
    for i in range(args.steps):
        step(i)
    e1.record(cur)
    _barrier(world, dev)
    ms = _max_over_ranks(e0.elapsed_time(e1), world, dev)
    val = B * world * args.steps / (ms / 1e3)
    out = {"metric": METRIC, "value": round(val, 1), "unit": "images/sec", "n_gpus": world, "steps": args.steps,
           "warmup": max(args.warmup, 3), "ms_per_step": round(ms / args.steps, 5), "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "impl": "standin-torch-graph",
           "config": {"model": MODEL, "global_batch": B * world, "parallelism": f"dp{world}",
                      "note": "stand-in, not the reference: PyTorch + cuDNN/cuBLAS (bf16 autocast) + NCCL all-reduce of one "
                              "flat bucket, whole step in one CUDA graph, H2D per step; the reference (TensorFlow) is not "
                              "installable here"},
           "clocks": sampler.stop(),
           "e2e": {"value": round(val, 1), "unit": "images/sec", "h2d_bytes_per_step": B * 784 * 4 + B * 8,
                   "d2h_bytes_per_step": 4, "final_loss": float(out_host[args.steps - 1])}}
    return out if rank == 0 else {}


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="tfdl", choices=["tfdl", "reference", "standin"])
    ap.add_argument("--batch-per-gpu", type=int, default=0, help="0 = the config's batch (64 CNN, 256 ResNet-50, 64 BERT)")
    ap.add_argument("--model", default="cnn", choices=["cnn", "resnet50", "bert"],
                    help="cnn = the reference model (BASELINE.json config 2, the headline); resnet50 / bert = configs 3 / 4")
    args = ap.parse_args()
    if args.model == "cnn" and not args.batch_per_gpu:
        args.batch_per_gpu = 64
    if args.model != "cnn" and args.steps == 2000:
        args.steps, args.warmup = 20, 5
    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:
            return  # launched like the other arm (torchrun for N > 1): one line, from rank 0
        print(json.dumps({"impl": "reference",
                          "unavailable": "the reference is a TensorFlow 2.x script: /root/reference has no setup.py/"
                                         "pyproject (pip install fails) and no tensorflow/keras/tfds wheel exists in "
                                         "/opt/wheelhouse or the image"}))
        return
    if args.model != "cnn":
        if args.impl == "standin":
            raise SystemExit("the ResNet-50 / BERT stand-ins live in tools/bench_models.py --impl standin")
        out = run_model(args)
    else:
        out = run_standin(args) if args.impl == "standin" else run_ours(args)
    if out:
        print(json.dumps(out), flush=True)
    try:  # leave the process group cleanly so the JSON line is the last line of output
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized():
            dist.destroy_process_group()
    except Exception:
        pass


if __name__ == "__main__":
    main()
