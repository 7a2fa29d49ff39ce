#!/usr/bin/env python
"""NVLink reference bandwidths on this box, measured with the library copy engine path (not a product path): peer copy
GPU0 -> GPU1 one direction and both directions at once, 1 GiB, CUDA events. These are the denominators the all-reduce
roofline fractions in profiles/ are reported against (next to the profiling recipe's 770 GB/s peer copy figure)."""
import json
import torch

n = 1 << 30
a0 = torch.empty(n, dtype=torch.uint8, device="cuda:0")
a1 = torch.empty(n, dtype=torch.uint8, device="cuda:1")
b0, b1 = torch.empty_like(a0), torch.empty_like(a1)
s0, s1 = torch.cuda.Stream("cuda:0"), torch.cuda.Stream("cuda:1")


def run(bidir):
    for _ in range(2):
        with torch.cuda.stream(s0):
            a1.copy_(a0, non_blocking=True)
        if bidir:
            with torch.cuda.stream(s1):
                b0.copy_(b1, non_blocking=True)
    torch.cuda.synchronize("cuda:0"); torch.cuda.synchronize("cuda:1")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s0):
        e0.record(s0)
        for _ in range(5):
            a1.copy_(a0, non_blocking=True)
        e1.record(s0)
    if bidir:
        with torch.cuda.stream(s1):
            for _ in range(5):
                b0.copy_(b1, non_blocking=True)
    torch.cuda.synchronize("cuda:0"); torch.cuda.synchronize("cuda:1")
    return 5 * n / (e0.elapsed_time(e1) * 1e-3) / 1e9


print("NVLINK_PEAKS " + json.dumps({"peer_copy_one_direction_GBps": run(False), "peer_copy_per_direction_when_bidirectional_GBps": run(True)}))
