"""cProfile of model.fit on the native engine (host-side fixed + per-step overhead). Run on a GPU box."""
import cProfile, io, os, pstats, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, tfdl
from tfdl.models import build_and_compile_cnn_model
K = int(sys.argv[1]) if len(sys.argv) > 1 else 200
s = tfdl.distribute.MirroredStrategy(devices=["gpu:0"])
with s.scope():
    m = build_and_compile_cnn_model(learning_rate=1e-3)
B = 64
xs = np.random.rand(256 * B, 28, 28, 1).astype(np.float32); ys = np.random.randint(0, 10, 256 * B).astype(np.int32)
ds = tfdl.data.Dataset.from_tensor_slices((xs, ys)).batch(B).repeat()
o = tfdl.data.Options(); o.experimental_distribute.auto_shard_policy = tfdl.data.AutoShardPolicy.OFF
ds = ds.with_options(o)
m.fit(ds, epochs=1, steps_per_epoch=20, verbose=0)
torch.cuda.synchronize()
for rep in range(3):
    t0 = time.perf_counter(); m.fit(ds, epochs=1, steps_per_epoch=K, verbose=0); torch.cuda.synchronize()
    print(f"fit({K}) wall {1e3 * (time.perf_counter() - t0):.2f} ms")
pr = cProfile.Profile(); pr.enable()
m.fit(ds, epochs=1, steps_per_epoch=K, verbose=0); torch.cuda.synchronize()
pr.disable()
st = io.StringIO(); pstats.Stats(pr, stream=st).sort_stats("cumulative").print_stats(28); print(st.getvalue()[:6000])
