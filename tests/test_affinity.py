"""CPU placement of a worker next to its GPU (utils/affinity.py): only ever narrows the process's CPU set, applies to
every thread, and leaves things alone when NVML has nothing to say, when the set would be too small for the workers that
share it, or when it is switched off."""
import os
import threading

import pytest
import torch

from tfdl.utils import affinity

pytestmark = pytest.mark.skipif(not hasattr(os, "sched_setaffinity"), reason="needs sched_setaffinity")


@pytest.fixture
def restore_affinity():
    before = os.sched_getaffinity(0)
    yield before
    for tid in os.listdir("/proc/self/task"):
        try:
            os.sched_setaffinity(int(tid), before)
        except OSError:
            pass


def test_bind_narrows_every_thread(monkeypatch, restore_affinity):
    allowed = sorted(restore_affinity)
    if len(allowed) < 4:
        pytest.skip("needs at least 4 usable CPUs")
    local = set(allowed[: len(allowed) // 2]) | {10_000}  # a CPU outside the allowed set is ignored
    monkeypatch.setattr(affinity, "device_cpus", lambda dev: local)
    monkeypatch.setenv("TFDL_CPU_AFFINITY", "1")
    stop = threading.Event()
    seen = {}
    t = threading.Thread(target=lambda: (stop.wait(5), seen.setdefault("aff", os.sched_getaffinity(0))))
    t.start()
    got = affinity.bind_to_device(torch.device("cuda", 0), local_world=1)
    stop.set()
    t.join()
    want = set(allowed[: len(allowed) // 2])
    assert got == want and os.sched_getaffinity(0) == want
    assert seen["aff"] == want  # a thread that existed before the call moved too


@pytest.mark.parametrize("case", ["off", "single_process_auto", "one_numa_node_auto", "no_nvml", "too_small",
                                  "already_local", "cpu_device"])
def test_bind_leaves_affinity_alone(monkeypatch, restore_affinity, case):
    allowed = set(restore_affinity)
    local = set(sorted(allowed)[:1])
    world = 8
    dev = torch.device("cuda", 0)
    monkeypatch.setenv("TFDL_CPU_AFFINITY", "auto")
    monkeypatch.setattr(affinity, "numa_nodes", lambda: 2)
    if case == "one_numa_node_auto":
        monkeypatch.setattr(affinity, "numa_nodes", lambda: 1)
        world, local = 2, set(sorted(allowed)[: max(1, len(allowed) // 2)])
    if case == "off":
        monkeypatch.setenv("TFDL_CPU_AFFINITY", "0")
    elif case == "single_process_auto":
        world, local = 1, set(sorted(allowed)[: max(1, len(allowed) // 2)])
    elif case == "no_nvml":
        local = None
    elif case == "already_local":
        local = allowed
    elif case == "cpu_device":
        dev = torch.device("cpu")
    monkeypatch.setattr(affinity, "device_cpus", lambda d: local)
    assert affinity.bind_to_device(dev, local_world=world) is None
    assert os.sched_getaffinity(0) == allowed


def test_device_cpus_without_a_gpu_is_none():
    assert affinity.device_cpus(torch.device("cuda", 0)) is None or torch.cuda.is_available()


def test_auto_binds_multi_process_launch_on_a_two_node_host(monkeypatch, restore_affinity):
    allowed = sorted(restore_affinity)
    if len(allowed) < 8:
        pytest.skip("needs at least 8 usable CPUs")
    local = set(allowed[: len(allowed) // 2])
    monkeypatch.setattr(affinity, "device_cpus", lambda dev: local)
    monkeypatch.setattr(affinity, "numa_nodes", lambda: 2)
    monkeypatch.setenv("TFDL_CPU_AFFINITY", "auto")
    assert affinity.bind_to_device(torch.device("cuda", 0), local_world=2) == local
    assert os.sched_getaffinity(0) == local


def test_numa_nodes_reads_the_host():
    assert affinity.numa_nodes() >= 1
