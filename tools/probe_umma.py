"""Run the hardware probes and print what the silicon does (see csrc/kernels/probe.cu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tfdl.ops import _lib
L = _lib.kernels(); p = _lib.ptr
torch.manual_seed(0)
for kc in (64, 32):
    rows, N = 256, 64
    A = torch.randn(rows, kc, device="cuda").bfloat16(); B = torch.randn(N, kc, device="cuda").bfloat16()
    for shift in (0, 8, 1, 3, 13, 100):
        for mode in (0, 1):
            D = torch.zeros(128, N, device="cuda")
            rc = L.tfdl_probe_shifted_umma(p(A), p(B), p(D), rows, N, kc, shift, mode, None)
            torch.cuda.synchronize()
            ref = A[shift:shift + 128].float() @ B.float().t()
            err = (D - ref).abs().max().item()
            print(f"kc={kc} shift={shift:3d} base_offset_mode={mode} rc={rc} max_err={err:.4f} {'OK' if err < 0.05 else 'WRONG'}")
rows, K = 64, 288
W = torch.randn(rows, K, device="cuda").bfloat16()
out = torch.zeros(rows * K, dtype=torch.bfloat16, device="cuda")
rc = L.tfdl_probe_tma_coremajor(p(W), p(out), rows, K, None)
torch.cuda.synchronize()
if rc != 0:
    print("tma_coremajor rc", rc, L.tfdl_last_error())
else:
    # expected: element (row, k) at ((row//8)*(K//8) + k//8)*64 + (row%8)*8 + k%8
    exp = W.view(rows // 8, 8, K // 8, 8).permute(0, 2, 1, 3).reshape(-1)
    print("tma_coremajor", "OK" if torch.equal(out, exp) else "WRONG", (out.float() - exp.float()).abs().max().item())
for krows in (64, 144):
    for lbo_mode in (0, 1):
        for shift in (0, 1, 13, 28):
            acols = 128 if lbo_mode == 0 else 64
            A = torch.randn(krows, acols, device="cuda").bfloat16()
            Bm = torch.randn(krows + 64, 32, device="cuda").bfloat16()
            D = torch.zeros(128, 32, device="cuda")
            rc = L.tfdl_probe_mn_umma(p(A), p(Bm), p(D), krows, shift, lbo_mode, None)
            torch.cuda.synchronize()
            Af = A.float() if lbo_mode == 0 else torch.cat([A.float(), A.float()], 1)
            ref = Af.t() @ Bm[shift:shift + krows].float()
            err = (D - ref).abs().max().item()
            err_lo = (D[:64] - ref[:64]).abs().max().item()
            print(f"mn-major krows={krows} lbo_mode={lbo_mode} shift={shift:2d} rc={rc} max_err={err:.4f} (rows<64: {err_lo:.4f}) {'OK' if err < 0.1 else 'WRONG'}")
