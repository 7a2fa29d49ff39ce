"""Multi-GPU checks, one process per GPU (launched by tests/test_multigpu.py through torch.distributed.run).
Prints one JSON line per rank-0 check; exits non-zero on the first failed assertion."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import tfdl  # noqa: E402
from tfdl import keras  # noqa: E402
from tfdl.ops.allreduce import P2PFusedAllReduce  # noqa: E402
from tfdl.ops.structs import AR_MULTIMEM, AR_NAMES, AR_ONESHOT, AR_PUSH, AR_PUSH2, AR_TWOSHOT  # noqa: E402


def main() -> None:
    import torch.distributed as dist

    strategy = tfdl.distribute.MirroredStrategy()
    rank, world = strategy.worker_index, strategy.num_workers
    dev = strategy.local_devices[0]
    ops = strategy.cross_device_ops
    assert isinstance(ops, P2PFusedAllReduce) and ops.world == world
    report = {"world": world}

    # ---- fused all-reduce + SGD, every algorithm, several sizes, back-to-back steps with skewed ranks ----
    for n in (1024, 225280, 4 * 1024 * 1024 + 64):
        params, grads = ops.alloc_arena(n, dev), ops.alloc_arena(n, dev)
        st = ops.fused_state(params, grads, momentum=True)
        algos = [AR_ONESHOT, AR_TWOSHOT, AR_PUSH2] + ([AR_MULTIMEM] if st.has_multicast else [])
        if n * 4 * (world - 1) <= 32 << 20:
            algos.append(AR_PUSH)
        report["multicast"] = bool(st.has_multicast)
        for algo in algos:
            torch.manual_seed(1)
            w0 = torch.randn(n, device=dev)
            params.copy_(w0)
            st.mom.zero_()
            for g in st.grads:
                g.tensor.zero_()
            st.set_hyper(0.1, 0.9, False, 1.0 / world)
            torch.cuda.synchronize(dev)
            dist.barrier()
            w_ref, v_ref = w0.clone(), torch.zeros(n, device=dev)
            steps = 9
            all_g = [[torch.Generator(device="cpu").manual_seed(100 * s + r) for r in range(world)] for s in range(steps)]
            for s in range(steps):
                par = s & 1
                gs = [torch.randn(n, generator=all_g[s][r]).to(dev) for r in range(world)]
                if (s + rank) % 2 == 0:
                    torch.cuda._sleep(20_000_000)  # skew the ranks: ~10 ms
                st.grads[par].tensor.copy_(gs[rank])
                # step counters that wrap (flag values are compared wrap-safe / for equality): every rank moves its
                # device epoch forward at the same point of its stream, in three hops of < 2^31 so that each looks
                # monotonic to the barrier; steps 5..8 then run with epochs 0xFFFFFFFE, 0xFFFFFFFF, 0, 1
                if s == 1:
                    st.counters[0:1].fill_(0x50000000)
                if s == 3:
                    st.counters[0:1].fill_(-0x60000000)  # 0xA0000000
                if s == 5:
                    st.counters[0:1].fill_(-3)
                st.step(par, algo)  # no host sync between steps
                gsum = gs[0].clone()
                for r in range(1, world):
                    gsum += gs[r]
                gsum *= 1.0 / world
                v_ref = 0.9 * v_ref - 0.1 * gsum
                w_ref = w_ref + v_ref
            torch.cuda.synchronize(dev)
            # every kernel also clears the OTHER parity's gradient buffer for the next step's backward pass
            assert float(st.grads[par ^ 1].tensor.abs().max()) == 0.0, f"other-parity gradients not cleared ({AR_NAMES[algo]})"
            if algo in (AR_ONESHOT, AR_PUSH, AR_PUSH2):  # rank-ordered fp32 sums: the host order exactly
                torch.testing.assert_close(params, w_ref, rtol=1e-5, atol=1e-5)
            else:
                torch.testing.assert_close(params, w_ref, rtol=1e-4, atol=1e-4)  # in-switch / sliced order differs
            torch.testing.assert_close(st.shadow.tensor.float(), params.bfloat16().float(), rtol=0, atol=0)
            # the sliced paths leave each rank with its own slice of the momentum; a checkpoint needs all of it
            ops.sync_optimizer_slots()
            torch.testing.assert_close(st.mom, v_ref, rtol=1e-4, atol=1e-4)
            gathered = [torch.empty_like(params) for _ in range(world)]
            dist.all_gather(gathered, params)
            for r in range(1, world):
                assert torch.equal(gathered[0], gathered[r]), f"rank {r} diverged ({AR_NAMES[algo]}, n={n})"
            report[f"fused_{AR_NAMES[algo]}_{n}"] = "ok"

    # ---- fused Adam (both slot arenas, bias correction from the kernel's own step counter) on every kind of path -------
    n = 225280
    params, grads = ops.alloc_arena(n, dev), ops.alloc_arena(n, dev)
    st = ops.fused_state(params, grads, momentum=False, adam=True)
    for algo in [AR_PUSH, AR_PUSH2, AR_TWOSHOT]:
        torch.manual_seed(2)
        w0 = torch.randn(n, device=dev)
        params.copy_(w0)
        st.mom.zero_()
        st.mom2.zero_()
        for g in st.grads:
            g.tensor.zero_()
        lr, b1, b2, eps = 1e-2, 0.9, 0.999, 1e-7
        torch.cuda.synchronize(dev)
        done = int(st.counters[0])  # steps this state has already run: Adam's t must restart at 1 for this algorithm
        st.set_hyper(lr, 0.0, False, 1.0 / world, b1, b2, eps, -float(done))
        torch.cuda.synchronize(dev)
        dist.barrier()
        w_ref, m_ref, v_ref = w0.clone(), torch.zeros(n, device=dev), torch.zeros(n, device=dev)
        for s in range(5):
            par = s & 1
            gs = [torch.randn(n, generator=torch.Generator(device="cpu").manual_seed(1000 * s + r)).to(dev) for r in range(world)]
            st.grads[par].tensor.copy_(gs[rank])
            st.step(par, algo)
            gsum = gs[0].clone()
            for r in range(1, world):
                gsum += gs[r]
            gsum *= 1.0 / world
            m_ref = b1 * m_ref + (1 - b1) * gsum
            v_ref = b2 * v_ref + (1 - b2) * gsum * gsum
            t = s + 1
            w_ref = w_ref - lr * (1 - b2 ** t) ** 0.5 / (1 - b1 ** t) * m_ref / (v_ref.sqrt() + eps)
        torch.cuda.synchronize(dev)
        torch.testing.assert_close(params, w_ref, rtol=2e-4, atol=2e-5)
        ops.sync_optimizer_slots()
        torch.testing.assert_close(st.mom, m_ref, rtol=1e-4, atol=1e-6)
        torch.testing.assert_close(st.mom2, v_ref, rtol=1e-4, atol=1e-7)
        gathered = [torch.empty_like(params) for _ in range(world)]
        dist.all_gather(gathered, params)
        for r in range(1, world):
            assert torch.equal(gathered[0], gathered[r]), f"rank {r} diverged (adam, {AR_NAMES[algo]})"
        report[f"fused_adam_{AR_NAMES[algo]}"] = "ok"

    # ---- plain in-place sum (odd sizes) --------------------------------------------------------------
    for n in (4, 7, 1001, (1 << 20) + 3):  # element counts that are not multiples of 4: the kernels finish the tail
        buf = ops.alloc_arena(n, dev)
        buf.copy_(torch.arange(n, device=dev, dtype=torch.float32) * (rank + 1))
        torch.cuda.synchronize(dev)
        dist.barrier()
        ops.allreduce_grads([buf])
        ops.allreduce_grads([buf])  # twice in a row: x -> x*T -> x*T*world
        torch.cuda.synchronize(dev)
        tot = sum(r + 1 for r in range(world))
        torch.testing.assert_close(buf, torch.arange(n, device=dev, dtype=torch.float32) * tot * world, rtol=1e-6, atol=0)
    report["plain_sum"] = "ok"

    # ---- metric totals (float64, bit-identical on every rank) and the chief's broadcast, both on the NVLink kernels ----
    for rep in range(3):  # three calls: both slot parities
        vals = np.array([rank + 1.0, 0.1 * (rank + 1) + rep, 1e15 + rank], dtype=np.float64)
        tot = ops.allreduce_scalars(vals)
        exp = np.array([sum(r + 1.0 for r in range(world)), sum(0.1 * (r + 1) + rep for r in range(world)),
                        sum(1e15 + r for r in range(world))])
        np.testing.assert_allclose(tot, exp, rtol=1e-15)
        allv = [torch.empty(3, dtype=torch.float64, device=dev) for _ in range(world)]
        dist.all_gather(allv, torch.from_numpy(tot).to(dev))
        assert all(torch.equal(allv[0], a) for a in allv)
    bc = ops.alloc_arena(4096 + 5, dev)
    bc.copy_(torch.arange(4101, device=dev, dtype=torch.float32) * (rank + 3))
    torch.cuda.synchronize(dev)
    ops.broadcast_params([bc])
    torch.testing.assert_close(bc, torch.arange(4101, device=dev, dtype=torch.float32) * 3, rtol=0, atol=0)
    report["scalars_and_broadcast"] = "ok"

    # ---- training: N GPUs at global batch 64*N == one GPU fed the same global batches -----------------------
    from tfdl.models import build_and_compile_cnn_model

    keras.clear_session()
    keras.layers.set_init_seed(7 + rank)  # different local inits: the chief's broadcast must win
    with strategy.scope():
        model = build_and_compile_cnn_model(learning_rate=0.05, momentum=0.9)
    rng = np.random.default_rng(0)
    G = 64 * world
    x, y = rng.random((G * 6, 28, 28, 1), dtype=np.float32), rng.integers(0, 10, size=G * 6)
    ds = tfdl.data.Dataset.from_tensor_slices((x, y)).batch(G)
    o = tfdl.data.Options()
    o.experimental_distribute.auto_shard_policy = tfdl.data.AutoShardPolicy.DATA
    w_init = model.get_weights()  # on rank 0 these are the values every replica starts from (chief broadcast)
    h = model.fit(ds.with_options(o), epochs=2, verbose=0)
    assert model.engine_name == "NativeCnnEngine"
    ws = [torch.from_numpy(w).to(dev) for w in model.get_weights()]
    for w in ws:
        got = [torch.empty_like(w) for _ in range(world)]
        dist.all_gather(got, w.contiguous())
        for r in range(1, world):
            assert torch.equal(got[0], got[r]), "mirrored variables diverged"
    report["train_identical"] = "ok"
    report["train_loss"] = h.history["loss"]
    report["algo"] = model._engine.algo_name
    if rank == 0:
        # single-GPU reference on the same global batches (per-replica 64*world is a multiple of 16)
        keras.clear_session()
        keras.layers.set_init_seed(7)
        solo = tfdl.distribute.MirroredStrategy(devices=[f"gpu:{dev.index}"])
        with solo.scope():
            ref = build_and_compile_cnn_model(learning_rate=0.05, momentum=0.9)
        hr = ref.fit(ds, epochs=2, verbose=0)
        np.testing.assert_allclose(h.history["loss"], hr.history["loss"], rtol=2e-2)
        for a, b, w0, v in zip(model.get_weights(), ref.get_weights(), w_init, ref.weights):
            # compare the accumulated UPDATES (bf16 activations + fp32 atomics: a few percent in the norm)
            ua, ub = (a - w0).astype(np.float64), (b - w0).astype(np.float64)
            rel = np.linalg.norm(ua - ub) / (np.linalg.norm(ub) + 1e-30)
            # bf16 activations, a different accumulation order per replica count and momentum over 12 steps:
            # 5-10 % in the update norm is what two correct runs differ by; a wrong reduction is off by O(1)
            assert rel < 0.2, (v.name, rel)
        report["train_matches_single_gpu"] = "ok"
        print("MGPU_REPORT " + json.dumps(report), flush=True)
    dist.barrier()
    strategy.close()


if __name__ == "__main__":
    main()
