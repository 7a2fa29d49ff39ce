"""Helpers to run N localhost "cluster nodes" in one test (the trick of README.md:61:
several nodes can be started on one machine because TF_CONFIG is per process)."""
import multiprocessing as mp
import os
import socket
import sys
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def free_ports(n):
    socks, ports = [], []
    for _ in range(n):
        s = socket.socket()
        s.bind(("127.0.0.1", 0))
        socks.append(s)
        ports.append(s.getsockname()[1])
    for s in socks:
        s.close()
    return ports


def _entry(fn, rank, addrs, q, args):
    try:
        if ROOT not in sys.path:
            sys.path.insert(0, ROOT)
        q.put((rank, "ok", fn(rank, addrs, *args)))
    except BaseException as e:  # noqa: BLE001
        q.put((rank, "err", f"{type(e).__name__}: {e}\n{traceback.format_exc()}"))


def run_workers(fn, world, args=(), timeout=120, expect_fail=()):
    """Run fn(rank, addrs, *args) in `world` spawned processes; returns results by rank."""
    ctx = mp.get_context("spawn")
    addrs = [f"127.0.0.1:{p}" for p in free_ports(world)]
    q = ctx.Queue()
    procs = [ctx.Process(target=_entry, args=(fn, r, addrs, q, args), daemon=True) for r in range(world)]
    for p in procs:
        p.start()
    out = {}
    try:
        for _ in range(world - len(expect_fail)):
            rank, status, val = q.get(timeout=timeout)
            out[rank] = (status, val)
    finally:
        for p in procs:
            p.join(timeout=10)
            if p.is_alive():
                p.terminate()
    return out
