"""Host runtime (libtfdl_runtime.so): crc32c, shuffle buffer, TCP group, ring all-reduce, failure detection."""
import os
import time

import numpy as np
import pytest

from _mp import free_ports, run_workers

native = pytest.importorskip("tfdl.runtime.native")
if not native.available():  # pragma: no cover
    pytest.skip("runtime library not built", allow_module_level=True)


def test_crc32c_known_vectors():
    assert native.crc32c(b"123456789") == 0xE3069283
    assert native.crc32c(b"") == 0
    assert native.crc32c(bytes(32)) == 0x8A9136AA
    # extend == one shot
    assert native.crc32c(b"6789", native.crc32c(b"12345")) == 0xE3069283
    from tensorboard.compat.tensorflow_stub.pywrap_tensorflow import masked_crc32c as tb_masked

    for blob in (b"hello", os.urandom(1000)):
        assert native.masked_crc32c(blob) == tb_masked(blob)


def test_shuffle_indices_is_permutation_with_buffer_locality():
    src = np.arange(5000, dtype=np.int64)
    out = native.shuffle_indices(src, 100, seed=7)
    assert sorted(out.tolist()) == src.tolist()
    # element emitted at position t can only come from the first t + buffer stream elements
    assert np.all(out[:4900] < np.arange(4900) + 100 + 1)
    assert not np.array_equal(out, src)
    assert np.array_equal(out, native.shuffle_indices(src, 100, seed=7))
    assert not np.array_equal(out, native.shuffle_indices(src, 100, seed=8))
    full = native.shuffle_indices(src, 10000, seed=1)
    assert sorted(full.tolist()) == src.tolist()


def test_gather_rows_and_fused_scale():
    rng = np.random.default_rng(0)
    x = rng.integers(0, 255, size=(1000, 28, 28, 1), dtype=np.uint8)
    idx = rng.integers(0, 1000, size=300)
    np.testing.assert_array_equal(native.gather_rows(x, idx), x[idx])
    out = np.empty((300, 28, 28, 1), np.float32)
    native.gather_rows_u8_to_f32(x, idx, 1 / 255.0, out.ctypes.data)
    np.testing.assert_allclose(out, x[idx].astype(np.float32) / 255, rtol=1e-6)
    with pytest.raises(Exception):
        native.gather_rows(x, np.array([1000]))


def _w_collectives(rank, addrs, n):
    from tfdl.runtime import native as nat

    g = nat.Group(rank, addrs, timeout_s=30)
    g.barrier()
    got = g.allgather_bytes(bytes([rank]) * 4)
    objs = g.allgather_object({"rank": rank, "pad": "x" * rank})
    a = np.arange(n, dtype=np.float32) * (rank + 1)
    g.allreduce_(a)
    d = np.full(3, rank + 0.5, dtype=np.float64)
    g.allreduce_(d)
    b = np.full(1000, float(rank), dtype=np.float32)
    g.broadcast_array(b)
    g.barrier()
    g.close()
    return got, [o["rank"] for o in objs], a, d, b


# the last two sizes give ring steps of 1 MB and more: the send half of each step runs on the helper thread (control.cc)
@pytest.mark.parametrize("world,n", [(2, 225034), (3, 1001), (4, 7), (2, 1_500_001), (3, 1_000_003)])
def test_group_collectives(world, n):
    res = run_workers(_w_collectives, world, args=(n,))
    assert all(v[0] == "ok" for v in res.values()), res
    tot = sum(r + 1 for r in range(world))
    for r in range(world):
        got, ranks, a, d, b = res[r][1]
        assert got == [bytes([k]) * 4 for k in range(world)]
        assert ranks == list(range(world))
        np.testing.assert_allclose(a, np.arange(n, dtype=np.float32) * tot, rtol=1e-6)
        np.testing.assert_allclose(d, sum(k + 0.5 for k in range(world)))
        assert np.all(b == 0.0)
    # bitwise identical on every rank
    for r in range(1, world):
        assert np.array_equal(res[0][1][2], res[r][1][2])


def _w_mismatch(rank, addrs):
    from tfdl.runtime import native as nat

    use = addrs if rank == 0 else addrs + ["127.0.0.1:1"]
    try:
        nat.Group(rank, use, timeout_s=3)
    except nat.CollectiveError as e:
        return str(e)
    return "no error"


def test_rendezvous_timeout_and_cluster_mismatch():
    t0 = time.time()
    with pytest.raises(native.CollectiveError, match="could not reach rank"):
        native.Group(0, ["127.0.0.1:1", "127.0.0.1:9"], timeout_s=0.5)
    assert time.time() - t0 < 10
    res = run_workers(_w_mismatch, 2)
    assert any("identical on every node" in v[1] or "could not reach" in v[1] or "timed out" in v[1]
               for v in res.values()), res


def _w_kill(rank, addrs):
    from tfdl.runtime import native as nat

    g = nat.Group(rank, addrs, timeout_s=20, hb_interval_s=0.05, hb_timeout_s=1.0)
    g.barrier()
    if rank == 1:
        os._exit(17)  # die mid-run without saying goodbye
    a = np.ones(1 << 20, dtype=np.float32)
    t0 = time.time()
    try:
        for _ in range(50):
            g.allreduce_(a)
    except nat.CollectiveError as e:
        return str(e), time.time() - t0, g.failed_rank
    return "no error", time.time() - t0, g.failed_rank


def test_killed_worker_is_reported_not_hung():
    res = run_workers(_w_kill, 3, timeout=60, expect_fail=(1,))
    for r in (0, 2):
        status, (msg, dt, failed) = res[r]
        assert status == "ok"
        assert "rank" in msg and msg != "no error", msg
        assert dt < 15
        assert failed in (1,), (r, failed, msg)


@pytest.mark.parametrize("ydt", [np.int32, np.int64])
def test_pack_records_matches_numpy(ydt):
    from tfdl.runtime import native as rt

    rng = np.random.default_rng(0)
    k, B = 5, 16
    xs = [rng.random((B, 28, 28, 1), dtype=np.float32) for _ in range(k)]
    ys = [rng.integers(0, 10, size=B).astype(ydt) for _ in range(k)]
    rec = B * 784 * 4 + B * 4
    ring = np.zeros((k + 1, rec), dtype=np.uint8)
    rt.pack_records(xs, ys, ring.ctypes.data, rec, threads=3)
    for i in range(k):
        np.testing.assert_array_equal(ring[i, :B * 784 * 4].view(np.float32), xs[i].reshape(-1))
        np.testing.assert_array_equal(ring[i, B * 784 * 4:].view(np.int32), ys[i].astype(np.int32))
    assert not ring[k].any()


def _w_desync(rank, addrs):
    """Workers that fell out of step: rank 0 reduces 10 float32 gradients while rank 1 already reduces 4 float64 metric
    totals (what happens when one worker's data runs out a step early)."""
    from tfdl.runtime import native

    g = native.Group(rank, addrs, timeout_s=10, hb_timeout_s=5)
    try:
        if rank == 0:
            g.allreduce_(np.ones(10, np.float32))
        else:
            g.allreduce_(np.ones(4, np.float64))
    except Exception as e:  # noqa: BLE001
        return f"{type(e).__name__}: {e}"
    finally:
        g.close()
    return "no error"


def test_ring_collectives_out_of_step_fail_with_a_message():
    res = run_workers(_w_desync, 2, timeout=60)
    assert all("out of step" in v[1] for v in res.values()), res


def _records_oracle(x, y, flat, k, B):
    rows = int(np.prod(x.shape[1:])) * x.itemsize
    out = np.zeros((k, B * rows + B * 4), dtype=np.uint8)
    for r in range(k):
        sel = flat[r * B:(r + 1) * B]
        out[r, :B * rows] = np.ascontiguousarray(x[sel]).view(np.uint8).reshape(-1)
        out[r, B * rows:] = y[sel].astype(np.int32).view(np.uint8)
    return out


@pytest.mark.parametrize("k,B,ydt", [(1, 64, np.int64), (7, 48, np.int32), (20, 64, np.int64), (3, 20, np.uint8)])
def test_gather_records_sync_and_async_match_numpy(k, B, ydt):
    """shuffle + batch + staging in one gather (tf_dist_example.py:31-33), in its blocking form and as the background
    job that fit() starts its steps behind: same bytes, records complete in order, nothing written past the k-th record."""
    from tfdl.runtime import native as rt

    rng = np.random.default_rng(k * 100 + B)
    x = rng.random((5000, 28, 28, 1), dtype=np.float32)
    y = rng.integers(0, 10, size=5000).astype(ydt)
    flat = rng.integers(0, 5000, size=k * B).astype(np.int64)
    want = _records_oracle(x, y, flat, k, B)
    rec = want.shape[1]
    ring = np.full((k + 1, rec), 0xAB, dtype=np.uint8)
    rt.gather_records(x, y, flat, k, B, ring.ctypes.data, rec, threads=8)
    np.testing.assert_array_equal(ring[:k], want)
    assert (ring[k] == 0xAB).all()
    for first in (1, min(2, k), k):
        ring[:] = 0xCD
        job = rt.GatherJob(x, y, flat.copy(), k, B, ring.ctypes.data, rec)
        job.wait(first)
        got = job.ready()
        assert first <= got <= k
        np.testing.assert_array_equal(ring[:got], want[:got])  # whatever is reported ready is complete
        job.wait(k)
        assert job.ready() == k
        job.end()
        np.testing.assert_array_equal(ring[:k], want)
        assert (ring[k] == 0xCD).all()


def test_gather_jobs_queue_behind_each_other_and_reject_bad_rows():
    from tfdl.runtime import native as rt

    rng = np.random.default_rng(5)
    x = rng.random((3000, 784), dtype=np.float32)
    y = rng.integers(0, 10, size=3000).astype(np.int64)
    k, B = 12, 64
    flats = [rng.integers(0, 3000, size=k * B).astype(np.int64) for _ in range(4)]
    rec = B * 784 * 4 + B * 4
    rings = [np.zeros((k, rec), dtype=np.uint8) for _ in flats]
    jobs = [rt.GatherJob(x, y, f, k, B, r.ctypes.data, rec) for f, r in zip(flats, rings)]  # four jobs in flight
    rt.gather_rows(x, flats[0][:300])  # a blocking call on the same pool while they run
    for j in reversed(jobs):
        j.end()
    for f, r in zip(flats, rings):
        np.testing.assert_array_equal(r, _records_oracle(x, y, f, k, B))
    bad = flats[0].copy()
    bad[17] = 3000
    with pytest.raises(RuntimeError, match="out of range"):
        rt.GatherJob(x, y, bad, k, B, rings[0].ctypes.data, rec)
    with pytest.raises(RuntimeError, match="out of range"):
        rt.gather_records(x, y, bad, k, B, rings[0].ctypes.data, rec)


def test_rendezvous_ignores_stray_connections():
    """Something else connects to a worker's port while the cluster is forming (a port scanner, a health probe): bytes that
    are not this protocol's hello, or a connection that closes half way through it, are dropped and the rendezvous goes on."""
    import socket
    import threading

    from tfdl.runtime import native

    addrs = [f"127.0.0.1:{p}" for p in free_ports(2)]
    groups, errors = {}, []

    def join(rank):
        try:
            groups[rank] = native.Group(rank, addrs, timeout_s=30, hb_timeout_s=5)
        except Exception as e:  # noqa: BLE001
            errors.append((rank, repr(e)))

    t0 = threading.Thread(target=join, args=(0,))
    t0.start()
    host, port = addrs[0].split(":")
    deadline = time.time() + 10
    while True:  # wait until rank 0 listens
        try:
            s = socket.create_connection((host, int(port)), timeout=1)
            break
        except OSError:
            assert time.time() < deadline
            time.sleep(0.05)
    s.sendall(os.urandom(64))  # not a hello
    s.close()
    s2 = socket.create_connection((host, int(port)), timeout=1)
    s2.sendall(b"\x01\x02\x03")  # closes in the middle of one
    s2.close()
    t1 = threading.Thread(target=join, args=(1,))
    t1.start()
    t0.join(40)
    t1.join(40)
    assert not errors and set(groups) == {0, 1}, errors
    try:
        out = {}

        def reduce(rank):
            a = np.full(1000, rank + 1.0, dtype=np.float32)
            groups[rank].allreduce_(a)
            out[rank] = a

        ts = [threading.Thread(target=reduce, args=(r,)) for r in (0, 1)]
        [t.start() for t in ts]
        [t.join(30) for t in ts]
        assert np.all(out[0] == 3.0) and np.all(out[1] == 3.0)
    finally:
        for g in groups.values():
            g.close()


def _w_big_ring_with_a_death(rank, addrs):
    """Large all-reduces (helper-thread path); rank 1 exits after the second one."""
    from tfdl.runtime import native as nat

    g = nat.Group(rank, addrs, timeout_s=20, hb_interval_s=0.1, hb_timeout_s=1.5)
    buf = np.ones(2_000_000, dtype=np.float32)
    t0 = time.time()
    try:
        for it in range(50):
            if rank == 1 and it == 2:
                os._exit(0)
            buf[:] = 1.0
            g.allreduce_(buf)
            assert buf[0] == 2.0 and buf[-1] == 2.0
    except nat.CollectiveError as e:
        return str(e), time.time() - t0, g.failed_rank
    finally:
        if rank == 0:
            g.close()
    return "no error", time.time() - t0, -1


def test_killed_worker_during_large_allreduce_is_reported_not_hung():
    res = run_workers(_w_big_ring_with_a_death, 2, timeout=90, expect_fail=(1,))
    status, out = res[0]
    assert status == "ok", res
    msg, dt, failed = out
    assert "rank 1" in msg and msg != "no error", msg
    assert dt < 30 and failed == 1
