#!/usr/bin/env python
"""Host CPU / GPU locality of this box: CPU count, the CPUs this process may use, NUMA nodes, and for every GPU NVML
sees the CPUs it reports as local (what `tfdl.utils.affinity.bind_to_device` pins a worker to).

    python tools/cpu_topology.py
"""
from __future__ import annotations

import glob
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def ranges(cpus) -> str:
    cpus = sorted(cpus)
    out, i = [], 0
    while i < len(cpus):
        j = i
        while j + 1 < len(cpus) and cpus[j + 1] == cpus[j] + 1:
            j += 1
        out.append(f"{cpus[i]}-{cpus[j]}" if j > i else str(cpus[i]))
        i = j + 1
    return ",".join(out)


def main() -> None:
    info = {"cpu_count": os.cpu_count(), "allowed": ranges(os.sched_getaffinity(0)), "numa_nodes": {}}
    for node in sorted(glob.glob("/sys/devices/system/node/node[0-9]*")):
        try:
            info["numa_nodes"][os.path.basename(node)] = open(os.path.join(node, "cpulist")).read().strip()
        except OSError:
            pass
    try:
        import pynvml

        pynvml.nvmlInit()
        gpus = {}
        for i in range(pynvml.nvmlDeviceGetCount()):
            h = pynvml.nvmlDeviceGetHandleByIndex(i)
            words = pynvml.nvmlDeviceGetCpuAffinity(h, ((os.cpu_count() or 1) + 63) // 64)
            cpus = {64 * w + b for w, m in enumerate(words) for b in range(64) if (int(m) >> b) & 1}
            pci = pynvml.nvmlDeviceGetPciInfo(h).busId
            gpus[f"gpu{i}"] = {"pci": pci.decode() if isinstance(pci, bytes) else pci, "local_cpus": ranges(cpus)}
        info["nvml"] = gpus
    except Exception as e:  # noqa: BLE001
        info["nvml"] = f"unavailable: {type(e).__name__}: {e}"
    try:
        import torch

        from tfdl.utils.affinity import device_cpus

        if torch.cuda.is_available():
            info["visible"] = {f"cuda:{i}": ranges(device_cpus(torch.device("cuda", i)) or []) for i in range(torch.cuda.device_count())}
    except Exception as e:  # noqa: BLE001
        info["visible"] = f"{type(e).__name__}: {e}"
    print("CPU_TOPOLOGY " + json.dumps(info))


if __name__ == "__main__":
    main()
