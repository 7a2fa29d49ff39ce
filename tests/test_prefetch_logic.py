"""The launch loop of the prefetched gather (keras/native_engine.py:_run_prefetched) without a GPU: the method runs on a
stand-in object whose run_pinned() records what it was asked to launch and checks that those records are already complete
in the ring at that moment - the property the device depends on (it reads the ring right after the launch)."""
import time
import types

import numpy as np
import pytest
import torch

from tfdl.keras.native_engine import NativeCnnEngine


class _Probe:
    def __init__(self, x, y, flat, k, B, slow=0.0):
        rows = int(np.prod(x.shape[1:])) * x.itemsize
        self.rec_bytes = B * rows + B * 4
        self.ring = torch.full((k + 2, self.rec_bytes), 0xEE, dtype=torch.uint8)
        self.host_out = torch.zeros((k + 2, 2))
        self.prof = {"gather": 0.0, "enqueue": 0.0}
        self.calls = []
        self.slow = slow
        self._want = np.zeros((k, self.rec_bytes), dtype=np.uint8)
        for r in range(k):
            sel = flat[r * B:(r + 1) * B]
            self._want[r, :B * rows] = np.ascontiguousarray(x[sel]).view(np.uint8).reshape(-1)
            self._want[r, B * rows:] = y[sel].astype(np.int32).view(np.uint8)

    def run_pinned(self, records, n_steps, host_out=None, first=0):
        assert records is self.ring and host_out is self.host_out
        got = self.ring[first:first + n_steps].numpy()
        np.testing.assert_array_equal(got, self._want[first:first + n_steps])  # complete BEFORE the launch
        self.calls.append((first, n_steps))
        if self.slow:
            time.sleep(self.slow)  # a slow launch: more records are ready by the next round


@pytest.mark.parametrize("k,B,slow", [(3, 16, 0.0), (5, 64, 0.0), (20, 64, 0.0), (20, 64, 0.002), (37, 32, 0.0005),
                                      (64, 64, 0.0)])
def test_every_record_is_launched_once_in_order_and_only_when_complete(k, B, slow):
    rng = np.random.default_rng(k * 1000 + B)
    x = rng.random((4000, 28, 28, 1), dtype=np.float32)
    y = rng.integers(0, 10, size=4000).astype(np.int64)
    flat = rng.integers(0, 4000, size=k * B).astype(np.int64)
    probe = _Probe(x, y, flat, k, B, slow)
    run = types.MethodType(NativeCnnEngine._run_prefetched, probe)
    run(x, y, flat, k, B, time.perf_counter())
    firsts = [c[0] for c in probe.calls]
    sizes = [c[1] for c in probe.calls]
    assert firsts[0] == 0 and sum(sizes) == k
    assert all(firsts[i + 1] == firsts[i] + sizes[i] for i in range(len(sizes) - 1))  # contiguous, in order
    assert sizes[0] in (2, 4) or sizes[0] == k  # the device is started on the first records, not on the whole chunk
    assert len(sizes) <= 2 + k // 4             # never a launch per step
    assert (probe.ring[k:] == 0xEE).all()       # nothing written past the chunk
    assert probe.prof["gather"] >= 0.0 and probe.prof["enqueue"] >= 0.0


def test_a_failing_launch_still_ends_the_gather_job():
    rng = np.random.default_rng(0)
    x = rng.random((500, 28, 28, 1), dtype=np.float32)
    y = rng.integers(0, 10, size=500).astype(np.int64)
    flat = rng.integers(0, 500, size=8 * 16).astype(np.int64)
    probe = _Probe(x, y, flat, 8, 16)

    def boom(*a, **k):
        raise RuntimeError("launch failed")

    probe.run_pinned = boom
    with pytest.raises(RuntimeError, match="launch failed"):
        types.MethodType(NativeCnnEngine._run_prefetched, probe)(x, y, flat, 8, 16, time.perf_counter())
    # the job was completed and freed in the finally block: the ring holds every record, and a new job can use the pool
    np.testing.assert_array_equal(probe.ring[:8].numpy(), probe._want)
