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


def test_tutorial_walkthrough_code_runs(tmp_path):
    """docs/TUTORIAL.md section 5 (the counterpart of the reference's walkthrough, README.md:70-156): its python blocks,
    concatenated, are a working training script - here on one CPU replica with two short epochs."""
    import re

    text = open(os.path.join(ROOT, "docs", "TUTORIAL.md")).read()
    section = text[text.index("## 5. Walk through"):text.index("## 6. Launching")]
    blocks = re.findall(r"```python\n(.*?)```", section, re.S)
    assert len(blocks) >= 7
    parts = []
    for b in blocks:
        if 'os.environ["TF_CONFIG"]' in b:  # the two-machine cluster of the reference: one local CPU replica instead
            b = "strategy = tfdl.distribute.MirroredStrategy(devices=['cpu'])\n"
        if "train_x" in b:
            continue
        parts.append(b.replace("epochs=10, steps_per_epoch=20", "epochs=2, steps_per_epoch=2")
                     .replace('"ckpt/', '"' + str(tmp_path) + "/"))
    script = tmp_path / "tutorial.py"
    script.write_text("import sys\nsys.path.insert(0, %r)\n" % ROOT + "\n".join(parts) +
                      "\nprint('HISTORY', history.history['loss'])\n")
    env = {k: v for k, v in os.environ.items() if k not in ("TF_CONFIG", "RANK", "WORLD_SIZE", "LOCAL_RANK")}
    env["CUDA_VISIBLE_DEVICES"] = ""
    r = subprocess.run([sys.executable, str(script)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                       timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-4000:]
    assert "HISTORY" in r.stdout
    assert any(f.startswith("ckpt-2") and f.endswith(".index") for f in os.listdir(tmp_path))


def test_example_cluster_with_an_evaluator_task(tmp_path):
    """chief + worker train over the TCP ring while an `evaluator` task of the same TF_CONFIG cluster (README.md:38-44,57
    of the reference) evaluates each checkpoint the chief writes."""
    import socket

    def port():
        with socket.socket() as s:
            s.bind(("127.0.0.1", 0))
            return s.getsockname()[1]

    cluster = {"chief": [f"127.0.0.1:{port()}"], "worker": [f"127.0.0.1:{port()}"], "evaluator": [f"127.0.0.1:{port()}"]}
    base = [sys.executable, os.path.join(ROOT, "examples", "mnist_multiworker.py"), "--device", "cpu", "--synthetic", "512",
            "--epochs", "2", "--steps-per-epoch", "2", "--communication", "RING", "--checkpoint-dir", str(tmp_path / "run")]
    env0 = {k: v for k, v in os.environ.items() if k not in ("TF_CONFIG", "RANK", "WORLD_SIZE", "LOCAL_RANK")}
    env0["CUDA_VISIBLE_DEVICES"] = ""
    procs = []
    for ttype, idx in (("evaluator", 0), ("chief", 0), ("worker", 0)):
        env = dict(env0, TF_CONFIG=json.dumps({"cluster": cluster, "task": {"type": ttype, "index": idx}}))
        procs.append((ttype, subprocess.Popen(base, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env)))
    outs = {}
    for ttype, p in procs:
        out, _ = p.communicate(timeout=600)
        assert p.returncode == 0, (ttype, out[-3000:])
        outs[ttype] = out
    evals = [json.loads(l) for l in outs["evaluator"].splitlines() if l.startswith("{") and "evaluator" in l]
    # it stops at the run's last checkpoint (step counter 4 = 2 epochs x 2 steps); an epoch that is overwritten between two
    # polls is skipped, as with TensorFlow's SidecarEvaluator
    assert 1 <= len(evals) <= 2 and evals[-1]["checkpoint"].endswith("ckpt-2") and evals[-1]["step"] == 4
    assert all("loss" in e and "sparse_categorical_accuracy" in e for e in evals)
    assert any(l.startswith("{") and '"replicas": 2' in l for l in outs["chief"].splitlines())  # two training replicas


def test_cpu_side_tools_run():
    """tools/fit_overhead.py (host cost of fit() on the index path, no GPU needed) and tools/cpu_topology.py."""
    env = {k: v for k, v in os.environ.items() if k not in ("TF_CONFIG", "RANK", "WORLD_SIZE", "LOCAL_RANK")}
    env["CUDA_VISIBLE_DEVICES"] = ""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "fit_overhead.py"), "--reps", "5"], stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-3000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["steps"] == 20 and 0 < out["fit_until_first_train_call_us"] <= out["fit_python_only_us"] < 1e5
    assert 0 < out["gather_job_first_2_records_us"] <= out["gather_job_all_records_us"]
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "cpu_topology.py")], stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout[-3000:]
    topo = json.loads([l for l in r.stdout.splitlines() if l.startswith("CPU_TOPOLOGY ")][-1][len("CPU_TOPOLOGY "):])
    assert topo["cpu_count"] >= 1 and topo["allowed"]


def test_api_parity_table_resolves():
    """docs/API_PARITY.md lists, for every symbol the reference uses, the name it has here: each of those
    names must resolve, and each implementation / test file it points to must exist."""
    import importlib
    import re

    import tfdl  # noqa: F401

    text = open(os.path.join(ROOT, "docs", "API_PARITY.md")).read()
    rows = [l for l in text.splitlines() if l.startswith("| `") or l.startswith("| chief") or l.startswith("| Keras")]
    assert len(rows) >= 26
    names = set()
    for row in rows:
        cells = [c.strip() for c in row.strip("|").split("|")]
        names.update(re.findall(r"`(tfdl\.[A-Za-z_.0-9]+)`", cells[1]))
        for path in re.findall(r"`([a-z_/]+\.(?:py|cu|cc))`", cells[2]):
            cands = [os.path.join(ROOT, "tensorflow-distributed-learning_b200", path), os.path.join(ROOT, path),
                     os.path.join(ROOT, "csrc", "kernels", path)]
            assert any(os.path.exists(c) for c in cands), path
        for t in re.findall(r"`(test_[a-z_]+\.py)`", cells[3]):
            assert os.path.exists(os.path.join(ROOT, "tests", t)), t
    assert len(names) >= 30
    for dotted in sorted(names):
        parts = dotted.split(".")
        obj = importlib.import_module(parts[0])
        for p in parts[1:]:
            if p.startswith("."):
                continue
            obj = getattr(obj, p)
        assert obj is not None, dotted
