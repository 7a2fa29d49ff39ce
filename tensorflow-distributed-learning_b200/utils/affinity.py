"""CPU placement of a worker process next to its GPU.

Under a one-process-per-GPU launch (torchrun, or TF_CONFIG workers with one GPU each - the deployment the reference
describes, /root/reference/README.md:21-29) every step-graph launch is a handful of MMIO writes from the process's CPU
to its GPU. On a two-socket host a process that the scheduler left on the other socket pays the inter-socket hop on
each of them, and the end-to-end rate of a synchronous job is that of its slowest rank. `bind_to_device` pins the
process (all of its threads, and therefore the input pipeline's worker pool created later) to the CPUs NVML reports as
local to the device.

By default this happens only for multi-process launches on hosts with more than one NUMA node.
`TFDL_CPU_AFFINITY=0` turns it off, `=1` forces it on (single process, single-node hosts). Nothing here is required for
correctness: every failure (no NVML, no topology information in a container, an empty intersection with the cgroup's
CPU set) leaves the affinity untouched.
"""
from __future__ import annotations

import os
from typing import Optional, Set

import torch


def device_cpus(device: torch.device) -> Optional[Set[int]]:
    """CPUs local to a CUDA device according to NVML, or None when that cannot be determined."""
    try:
        import pynvml

        pynvml.nvmlInit()
        props = torch.cuda.get_device_properties(device)
        uuid = str(getattr(props, "uuid", "") or "")
        handle = None
        if uuid:
            try:
                handle = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
            except Exception:  # noqa: BLE001
                handle = None
        if handle is None:
            bus = f"{getattr(props, 'pci_domain_id', 0):08x}:{props.pci_bus_id:02x}:{props.pci_device_id:02x}.0"
            handle = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        n_cpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(handle, (n_cpu + 63) // 64)
        cpus = {64 * w + b for w, mask in enumerate(words) for b in range(64) if (int(mask) >> b) & 1}
        return cpus or None
    except Exception:  # noqa: BLE001
        return None


def numa_nodes() -> int:
    """Number of memory nodes the kernel exposes (1 when that cannot be read)."""
    try:
        return max(1, sum(1 for d in os.listdir("/sys/devices/system/node") if d.startswith("node") and d[4:].isdigit()))
    except OSError:
        return 1


def bind_to_device(device: torch.device, local_world: int = 1) -> Optional[Set[int]]:
    """Pin every thread of this process to the CPUs local to `device`. Returns the CPU set that was applied, or None
    when nothing was changed. Only narrows: the result is the intersection with what the process may already use, and a
    set too small for the workers that share it (fewer than 4 CPUs for each of the local_world / 2 workers of a socket) is not
    applied."""
    mode = os.environ.get("TFDL_CPU_AFFINITY", "auto")
    if mode == "0" or device.type != "cuda" or not hasattr(os, "sched_setaffinity"):
        return None
    if mode != "1" and (local_world <= 1 or numa_nodes() <= 1):
        return None  # a single process, or a host with one memory node (most VMs): nothing to gain from pinning
    local = device_cpus(device)
    if not local:
        return None
    try:
        allowed = os.sched_getaffinity(0)
        target = allowed & local
        if not target or target == allowed or len(target) < 2 * max(1, local_world):
            return None
        tids = [int(t) for t in os.listdir("/proc/self/task")] if os.path.isdir("/proc/self/task") else [0]
        for tid in tids:
            try:
                os.sched_setaffinity(tid, target)
            except OSError:
                pass  # a thread that exited meanwhile
        os.sched_setaffinity(0, target)
        return set(target)
    except OSError:
        return None
