#!/bin/bash
# ncu --set full of the transformer / ResNet support kernels (attention, cross-entropy, LayerNorm, BatchNorm), one launch
# each, driven by a small script at BERT / ResNet layer shapes. Plain run first. Output: gpurun_out/model_kernels.ncu-rep
set -u
mkdir -p gpurun_out
cat > /tmp/mk.py <<'PY'
import sys, torch
sys.path.insert(0, ".")
from tfdl.ops.attention import attention
from tfdl.ops.xent import sparse_softmax_xent
from tfdl.ops.layernorm import layer_norm
from tfdl.ops.batchnorm import batch_norm_train
torch.manual_seed(0)
for it in range(2):
    qkv = torch.randn(64, 128, 2304, device="cuda").bfloat16().requires_grad_()
    attention(qkv, 12).float().sum().backward()
    buf = torch.randn(8192, 30528, device="cuda").bfloat16(); logits = buf[:, :30522].detach().requires_grad_()
    sparse_softmax_xent(logits, torch.randint(0, 30522, (8192,), device="cuda"), -100).sum().backward()
    x = torch.randn(8192, 768, device="cuda").bfloat16().requires_grad_()
    g = torch.ones(768, device="cuda", requires_grad=True); b = torch.zeros(768, device="cuda", requires_grad=True)
    layer_norm(x, g, b, 1e-12, x.detach()).float().sum().backward()
    xb = torch.randn(256, 56, 56, 256, device="cuda").bfloat16().requires_grad_()
    g2 = torch.ones(256, device="cuda", requires_grad=True); b2 = torch.zeros(256, device="cuda", requires_grad=True)
    batch_norm_train(xb, g2, b2, torch.zeros(256, device="cuda"), torch.ones(256, device="cuda"), 0.01, 1e-3, True, xb.detach()).float().sum().backward()
torch.cuda.synchronize(); print("ok")
PY
python /tmp/mk.py > gpurun_out/mk_plain.log 2>&1 || { echo plain failed; tail -3 gpurun_out/mk_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:'attention_|softmax_xent|layernorm_|bn_' -s 11 -c 11 -f \
    -o gpurun_out/model_kernels python /tmp/mk.py > gpurun_out/ncu_mk.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/model_kernels.ncu-rep
