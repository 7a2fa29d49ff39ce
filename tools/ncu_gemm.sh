#!/bin/bash
# ncu --set full capture of the CTA-pair TMA GEMM (one launch, 8192 x 3072 x 768 and 4096^3), after a plain run exited 0.
set -u
mkdir -p gpurun_out
cat > /tmp/gemm_once.py <<'PY'
import sys, torch
sys.path.insert(0, ".")
from tfdl.ops.gemm import gemm_tma
for M, N, K in [(8192, 3072, 768), (4096, 4096, 4096)]:
    a = torch.randn(M, K, device="cuda").bfloat16(); b = torch.randn(N, K, device="cuda").bfloat16()
    for _ in range(3):
        out = gemm_tma(a, b)
torch.cuda.synchronize(); print("ok")
PY
python /tmp/gemm_once.py > gpurun_out/gemm_plain.log 2>&1 || { echo plain failed; tail -3 gpurun_out/gemm_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:tma_gemm_pair -s 2 -c 1 -f -o gpurun_out/gemm_pair_a python /tmp/gemm_once.py > gpurun_out/ncu_gemm_a.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:tma_gemm_pair -s 5 -c 1 -f -o gpurun_out/gemm_pair_b python /tmp/gemm_once.py > gpurun_out/ncu_gemm_b.log 2>&1
ls -la gpurun_out/gemm_pair_*.ncu-rep
