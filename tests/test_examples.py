"""The example that mirrors the reference script runs as a 2-task local cluster (README.md:61 of the reference)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_mnist_multiworker_example_two_cpu_workers(tmp_path):
    cmd = [sys.executable, os.path.join(ROOT, "examples", "launch_local_cluster.py"), "--workers", "2", "--",
           sys.executable, os.path.join(ROOT, "examples", "mnist_multiworker.py"), "--device", "cpu", "--synthetic", "1024",
           "--epochs", "2", "--steps-per-epoch", "3", "--communication", "RING", "--checkpoint-dir", str(tmp_path / "run")]
    env = {k: v for k, v in os.environ.items() if k not in ("TF_CONFIG", "RANK", "WORLD_SIZE", "LOCAL_RANK")}
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-4000:]
    reports = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{") and "replicas" in l]
    assert len(reports) == 1 and reports[0]["replicas"] == 2  # only the chief reports
    files = os.listdir(tmp_path / "run")
    assert any(f.startswith("ckpt-2") and f.endswith(".index") for f in files), files  # chief-only checkpoints
    assert os.path.isdir(tmp_path / "run" / "logs")
