"""Fused NVLink all-reduce + SGD and 2..8-GPU training: spawns one process per GPU (torch.distributed.run)."""
import json
import os
import subprocess
import sys

import pytest
import torch

from _mp import free_ports

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
HERE = os.path.dirname(os.path.abspath(__file__))


def _run(n):
    port = free_ports(1)[0]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(HERE, "mgpu_worker.py")]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    out_dir = os.path.join(os.path.dirname(HERE), "gpurun_out")
    if os.path.isdir(out_dir):  # keep the workers' output where a remote run can be read back
        with open(os.path.join(out_dir, f"mgpu_worker_N{n}.log"), "w") as f:
            f.write(r.stdout)
    assert r.returncode == 0, r.stdout[-6000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("MGPU_REPORT ")]
    assert line, r.stdout[-3000:]
    return json.loads(line[-1][len("MGPU_REPORT "):])


def test_all_visible_gpus():
    n = torch.cuda.device_count()
    rep = _run(n)
    assert rep["world"] == n and rep["train_identical"] == "ok" and rep["train_matches_single_gpu"] == "ok"
    assert rep["plain_sum"] == "ok" and rep.get("scalars_and_broadcast", "ok" if n == 1 else None) == "ok"
    assert all(v == "ok" for k, v in rep.items() if k.startswith("fused_"))


def test_two_gpus_when_more_are_visible():
    if torch.cuda.device_count() <= 2:
        pytest.skip("covered by test_all_visible_gpus")
    assert _run(2)["train_matches_single_gpu"] == "ok"
