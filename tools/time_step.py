"""Whole-step timing variants: (a) the executor's per-step graph launches with device-resident inputs,
(b) the same without the per-step input copies, to separate graph-launch overhead from kernel time."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, tfdl
from tfdl.models import build_and_compile_cnn_model
s = tfdl.distribute.MirroredStrategy(devices=["gpu:0"])
with s.scope():
    m = build_and_compile_cnn_model(1e-3)
e = m._make_engine(); e.prepare(64)
lib = e.lib
rec = e.rec_bytes
pool = torch.randint(0, 255, (512, rec), dtype=torch.uint8, device="cuda")
pool[:, 64*784*4:] = 0
def t(fn, n):
    fn(50); e.stream.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(e.stream); fn(n); e1.record(e.stream); e.stream.synchronize()
    return 1000 * e0.elapsed_time(e1) / n
print("run_device (2 D2D copies + graph launch per step): %.2f us/step" % t(lambda n: e.run_device(pool, n), 2000))
def steps_only(n):
    for _ in range(n):
        lib.tfdl_cnn_program_step(e._prog, C.c_void_p(e.stream.cuda_stream))
print("graph launch only (inputs left in place):           %.2f us/step" % t(steps_only, 2000))
