#!/usr/bin/env python
"""Throughput of the larger BASELINE.json configs (ResNet-50 bf16 256/GPU; BERT-base MLM bf16 seq 128):
this framework (TorchEngine + tcgen05 Dense GEMMs + fused NVLink all-reduce/SGD) vs the stand-in
(eager PyTorch + DDP/NCCL, same model math). Device-resident synthetic data, CUDA-event timing, max over ranks.

    python tools/bench_models.py --model resnet50 --batch 256 --steps 20
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/bench_models.py --model bert --batch 32
"""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch


def timed(step, steps, warmup, world, dev):
    import torch.distributed as dist
    for _ in range(warmup):
        step()
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize(dev)
    t = torch.tensor([e0.elapsed_time(e1) / steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet50", choices=["resnet50", "bert"])
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="both", choices=["tfdl", "standin", "both"])
    ap.add_argument("--profile", action="store_true", help="after timing: torch.profiler kernel table of 2 steps (diagnosis only)")
    a = ap.parse_args()
    world, rank, lrank = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
    dev = torch.device("cuda", lrank)
    torch.cuda.set_device(dev)
    out = {"model": a.model, "per_gpu_batch": a.batch, "n_gpus": world}
    g = torch.Generator(device="cpu").manual_seed(rank)
    if a.model == "resnet50":
        x = torch.rand((a.batch, 224, 224, 3), generator=g).to(dev)
        y = torch.randint(0, 1000, (a.batch,), generator=g).to(dev)
        unit = "images/sec"
    else:
        x = torch.randint(0, 30522, (a.batch, 128), generator=g).to(dev)
        y = torch.where(torch.rand((a.batch, 128), generator=g) < 0.15, x.cpu(), torch.full_like(x.cpu(), -100)).to(dev)
        unit = "sequences/sec"
    if a.impl in ("tfdl", "both"):
        import tfdl
        from tfdl import keras
        from tfdl.parallel.values import PerReplica
        keras.mixed_precision.set_global_policy("mixed_bfloat16")
        strategy = tfdl.distribute.MirroredStrategy() if world > 1 else tfdl.distribute.MirroredStrategy(devices=[f"gpu:{lrank}"])
        with strategy.scope():
            if a.model == "resnet50":
                from tfdl.models import build_and_compile_resnet50
                model = build_and_compile_resnet50(0.1, 0.9)
            else:
                from tfdl.models import build_and_compile_bert_mlm
                model = build_and_compile_bert_mlm(1e-3, optimizer="sgd")
        eng = model._make_engine()
        batch = PerReplica([(x, y)])
        ms = timed(lambda: eng.train_step(batch), a.steps, a.warmup, world, dev)
        out["tfdl"] = {"ms_per_step": ms, "value": a.batch * world / ms * 1e3, "unit": unit, "engine": type(eng).__name__,
                       "fused_allreduce_sgd": eng._fused is not None,
                       "allreduce": getattr(getattr(eng, "_fused", None), "torch_algo", None), "params": model.count_params()}
        if a.profile and rank == 0:
            from torch.profiler import ProfilerActivity, profile
            with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
                for _ in range(2):
                    eng.train_step(batch)
                torch.cuda.synchronize()
            print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=32, max_name_column_width=70))
        del model, eng
        torch.cuda.empty_cache()
    if a.impl in ("standin", "both"):
        import torch.nn as nn, torch.nn.functional as F
        if world > 1:
            import torch.distributed as dist
            if not dist.is_initialized():
                dist.init_process_group("nccl", device_id=dev)
        if a.model == "resnet50":
            class Bott(nn.Module):
                def __init__(s, cin, f, stride, proj):
                    super().__init__()
                    s.c1, s.b1 = nn.Conv2d(cin, f, 1, stride), nn.BatchNorm2d(f)
                    s.c2, s.b2 = nn.Conv2d(f, f, 3, 1, 1), nn.BatchNorm2d(f)
                    s.c3, s.b3 = nn.Conv2d(f, 4 * f, 1), nn.BatchNorm2d(4 * f)
                    s.p = nn.Sequential(nn.Conv2d(cin, 4 * f, 1, stride), nn.BatchNorm2d(4 * f)) if proj else None
                def forward(s, x):
                    sc = s.p(x) if s.p is not None else x
                    y = F.relu(s.b1(s.c1(x))); y = F.relu(s.b2(s.c2(y))); y = s.b3(s.c3(y))
                    return F.relu(y + sc)
            layers, cin = [nn.Conv2d(3, 64, 7, 2, 3), nn.BatchNorm2d(64), nn.ReLU(), nn.MaxPool2d(3, 2, 1)], 64
            for f, n, st in [(64, 3, 1), (128, 4, 2), (256, 6, 2), (512, 3, 2)]:
                for b in range(n):
                    layers.append(Bott(cin, f, st if b == 0 else 1, b == 0)); cin = 4 * f
            net = nn.Sequential(*layers, nn.AdaptiveAvgPool2d(1), nn.Flatten(), nn.Linear(2048, 1000)).to(dev).to(memory_format=torch.channels_last)
            xin = x.permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last)
            lossf = lambda m: F.cross_entropy(m(xin).float(), y)
        else:
            from transformers import BertConfig, BertForMaskedLM
            net = BertForMaskedLM(BertConfig(attn_implementation="sdpa")).to(dev)
            lossf = lambda m: m(input_ids=x, labels=y).loss
        model = nn.parallel.DistributedDataParallel(net, device_ids=[lrank]) if world > 1 else net
        opt = torch.optim.SGD(model.parameters(), lr=1e-3, momentum=0.9)
        def step():
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = lossf(model)
            opt.zero_grad(set_to_none=True); loss.backward(); opt.step()
        ms = timed(step, a.steps, a.warmup, world, dev)
        out["standin_torch_ddp"] = {"ms_per_step": ms, "value": a.batch * world / ms * 1e3, "unit": unit}
    if rank == 0:
        print("MODEL_BENCH " + json.dumps(out))


if __name__ == "__main__":
    main()
