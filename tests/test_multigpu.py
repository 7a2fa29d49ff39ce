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


def test_tf_config_workers_one_gpu_each():
    """The reference's launch form: one process per cluster task with its own TF_CONFIG (README.md:156-162)."""
    n = min(torch.cuda.device_count(), 2)
    if n < 2:
        pytest.skip("needs 2 GPUs")
    ports = free_ports(2)
    cluster = {"worker": [f"127.0.0.1:{p}" for p in ports]}
    procs = []
    for i in range(n):
        env = dict(os.environ, TF_CONFIG=json.dumps({"cluster": cluster, "task": {"type": "worker", "index": i}}))
        for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
            env.pop(k, None)
        procs.append(subprocess.Popen([sys.executable, os.path.join(HERE, "tfconfig_gpu_worker.py")], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=600)[0] for p in procs]
    reps = []
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o[-4000:]
        line = [l for l in o.splitlines() if l.startswith("TFCFG_REPORT ")]
        assert line, o[-2000:]
        reps.append(json.loads(line[-1][len("TFCFG_REPORT "):]))
    assert all(r["replicas"] == n and r["engine"] == "NativeCnnEngine" and r["ops"] == "P2PFusedAllReduce" for r in reps), reps
    assert reps[0]["wsum"] == reps[1]["wsum"] and reps[0]["wabs"] == reps[1]["wabs"]  # mirrored variables stay identical
    assert reps[0]["loss"] == reps[1]["loss"]


def test_killed_gpu_worker():
    """A GPU worker dies mid-run (README.md:65-68 only describes the happy path): the survivor's collective
    kernel is released by the heartbeat's abort word and fit() raises, naming the dead rank, within seconds."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    ports = free_ports(2)
    cluster = {"worker": [f"127.0.0.1:{p}" for p in ports]}
    procs = []
    for i in range(2):
        env = dict(os.environ, TF_CONFIG=json.dumps({"cluster": cluster, "task": {"type": "worker", "index": i}}))
        for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
            env.pop(k, None)
        procs.append(subprocess.Popen([sys.executable, os.path.join(HERE, "gpu_fault_worker.py")], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    for p in procs:
        try:
            outs.append(p.communicate(timeout=180)[0])
        except subprocess.TimeoutExpired:
            p.kill()
            outs.append(p.communicate()[0] + "\nTIMEOUT")
    assert procs[1].returncode == 43, outs[1][-2000:]  # the injected exit
    line = [l for l in outs[0].splitlines() if l.startswith("FAULT_REPORT ")]
    assert line, outs[0][-4000:]
    rep = json.loads(line[-1][len("FAULT_REPORT "):])
    assert rep["engine"] == "NativeCnnEngine"
    assert rep["error"] and "rank 1" in rep["error"], rep
    assert rep["seconds"] < 60, rep


def test_single_process_mirrored_strategy():
    """MirroredStrategy() in a plain python process (no torchrun, no TF_CONFIG) over every visible GPU."""
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs 2 GPUs")
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT")}
    r = subprocess.run([sys.executable, os.path.join(HERE, "mirrored_single_process_worker.py")], env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-6000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("MIRRORED_REPORT ")]
    assert line, r.stdout[-3000:]
    rep = json.loads(line[-1][len("MIRRORED_REPORT "):])
    assert rep["gpus"] == n and rep["engine"] == "MirroredNativeEngine" and rep["ops"] == "MirroredPeerAllReduce", rep
    assert rep["replicas_identical"] and rep["momentum_gathered"], rep
