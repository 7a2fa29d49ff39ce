#!/usr/bin/env python
"""Gradient all-reduce bus-bandwidth sweep (BASELINE.json config 5): this framework's NVLink kernels
vs NCCL (the stand-in for the reference's TF NCCL CollectiveOps), 1 KB - 1 GB, launched one process
per GPU:

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/allreduce_sweep.py [--max-mb 1024]

Each measurement is ITERS back-to-back collectives captured in one CUDA graph (no host in the loop),
timed with CUDA events, max over ranks. busbw = (S / t) * 2 (N-1) / N (nccl-tests convention).
"fused" = all-reduce + 1/N scale + SGD update + bf16 shadow refresh in one kernel (what training uses);
"sum"   = plain in-place sum.
Rank 0 writes profiles/allreduce_sweep_N<N>.json and prints a markdown table."""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import tfdl  # noqa: E402
from tfdl.ops.structs import AR_MULTIMEM, AR_NAMES, AR_ONESHOT, AR_PUSH, AR_PUSH2, AR_TWOSHOT  # noqa: E402


def timed_graph(fn, stream, iters, reps, dist, dev):
    with torch.cuda.stream(stream):
        fn()
        stream.synchronize()
        dist.barrier()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=stream):
            for _ in range(iters):
                fn()
        g.replay()
        stream.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps):
            g.replay()
        e1.record(stream)
    stream.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / (iters * reps)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()) * 1e-3  # seconds per collective


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=int, default=1024)
    ap.add_argument("--min-kb", type=int, default=1)
    ap.add_argument("--algos", default="oneshot,twoshot,multimem,push,push2")
    ap.add_argument("--tag", default="")
    ap.add_argument("--no-sum", action="store_true")
    args = ap.parse_args()
    import torch.distributed as dist

    strategy = tfdl.distribute.MirroredStrategy()
    ops = strategy.cross_device_ops
    rank, world, dev = ops.rank, ops.world, strategy.local_devices[0]
    stream = torch.cuda.Stream(dev)
    rows = []
    size = args.min_kb * 1024
    while size <= args.max_mb * 1024 * 1024:
        n = size // 4
        iters = 20 if size <= (64 << 20) else 5
        reps = 5 if size <= (64 << 20) else 2
        params, grads = ops.alloc_arena(n, dev), ops.alloc_arena(n, dev)
        st = ops.fused_state(params, grads, momentum=False)
        st.set_hyper(1e-3, 0.0, False, 1.0 / world)
        row = {"bytes": size}
        algos = [AR_TWOSHOT] + ([AR_MULTIMEM] if st.has_multicast else [])
        if size * (world - 1) <= (32 << 20):
            algos = [AR_ONESHOT, AR_PUSH] + algos
        if size <= (64 << 20):
            algos.append(AR_PUSH2)
        algos = [a for a in algos if AR_NAMES[a] in args.algos.split(",")]
        for algo in algos:
            par = [0]

            def fused():
                st.step(par[0], algo)
                par[0] ^= 1

            with torch.cuda.stream(stream):
                t = timed_graph(fused, stream, iters, reps, dist, dev)
            if rank == 0:
                print(f"  size {size} fused {AR_NAMES[algo]} ok {t * 1e6:.2f} us", flush=True)
            row[f"fused_{AR_NAMES[algo]}_us"] = t * 1e6
            row[f"fused_{AR_NAMES[algo]}_busbw"] = size / t * 2 * (world - 1) / world / 1e9
        for algo in [a for a in algos if a in (AR_TWOSHOT, AR_MULTIMEM) and not args.no_sum]:
            plain = ops._plain.get(grads.data_ptr())
            if plain is None:
                # one state = one stream: create (and first use) it on the stream that will run it
                with torch.cuda.stream(stream):
                    ops.allreduce_grads([grads])
                    stream.synchronize()
                plain = ops._plain[grads.data_ptr()]
            with torch.cuda.stream(stream):
                t = timed_graph(lambda: plain.allreduce_sum(algo), stream, iters, reps, dist, dev)
            if rank == 0:
                print(f"  size {size} sum {AR_NAMES[algo]} ok {t * 1e6:.2f} us", flush=True)
            row[f"sum_{AR_NAMES[algo]}_us"] = t * 1e6
            row[f"sum_{AR_NAMES[algo]}_busbw"] = size / t * 2 * (world - 1) / world / 1e9
        buf = torch.zeros(n, device=dev)
        with torch.cuda.stream(stream):
            t = timed_graph(lambda: dist.all_reduce(buf), stream, iters, reps, dist, dev)
        row["nccl_us"] = t * 1e6
        row["nccl_busbw"] = size / t * 2 * (world - 1) / world / 1e9
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
        del params, grads, st, buf
        ops._arenas.clear()
        ops._states.clear()
        ops._plain.clear()
        torch.cuda.empty_cache()
        size *= 4 if size < (1 << 20) else 2
    if rank == 0:
        os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
        out = os.path.join(ROOT, "gpurun_out", f"allreduce_sweep_N{world}{args.tag}.json")
        os.makedirs(os.path.dirname(out), exist_ok=True)
        json.dump({"world": world, "rows": rows}, open(out, "w"), indent=1)
        keys = [k for k in rows[-1] if k.endswith("_busbw")]
        print("| bytes | " + " | ".join(k.replace("_busbw", "") + " GB/s" for k in keys) + " | best fused us | nccl us |")
        print("|---|" + "---|" * (len(keys) + 2))
        for r in rows:
            best = min(v for k, v in r.items() if k.startswith("fused_") and k.endswith("_us"))
            print(f"| {r['bytes']} | " + " | ".join(f"{r.get(k, float('nan')):.1f}" for k in keys) +
                  f" | {best:.1f} | {r['nccl_us']:.1f} |")
    dist.barrier()
    strategy.close()


if __name__ == "__main__":
    main()
