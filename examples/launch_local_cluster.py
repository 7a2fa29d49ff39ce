#!/usr/bin/env python
"""Start N cluster tasks of a training script on THIS machine, each with its own TF_CONFIG (the reference README's
"several nodes on one physical machine", README.md:61). Everything after `--` is the command to run per task.

  python examples/launch_local_cluster.py --workers 2 -- python examples/mnist_multiworker.py --device cpu --synthetic 4096
  python examples/launch_local_cluster.py --workers 2 --gpus -- python examples/mnist_multiworker.py
"""
from __future__ import annotations

import argparse
import json
import os
import socket
import subprocess
import sys


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def main() -> int:
    if "--" not in sys.argv:
        print(__doc__)
        return 2
    split = sys.argv.index("--")
    ap = argparse.ArgumentParser()
    ap.add_argument("--workers", type=int, default=2)
    ap.add_argument("--chief", action="store_true", help="task 0 is a dedicated 'chief' job instead of worker 0")
    ap.add_argument("--gpus", action="store_true", help="pin task i to GPU i (adds --device gpu:i)")
    a = ap.parse_args(sys.argv[1:split])
    cmd = sys.argv[split + 1:]
    addrs = [f"127.0.0.1:{free_port()}" for _ in range(a.workers)]
    cluster = {"chief": addrs[:1], "worker": addrs[1:]} if a.chief else {"worker": addrs}
    procs = []
    for i in range(a.workers):
        task = {"type": "chief", "index": 0} if (a.chief and i == 0) else {"type": "worker", "index": i - (1 if a.chief else 0)}
        env = dict(os.environ, TF_CONFIG=json.dumps({"cluster": cluster, "task": task}))
        procs.append(subprocess.Popen(cmd + (["--device", f"gpu:{i}"] if a.gpus else []), env=env))
    rc = 0
    for p in procs:
        rc = rc or p.wait()
    return rc


if __name__ == "__main__":
    sys.exit(main())
