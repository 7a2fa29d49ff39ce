#!/bin/bash
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_multigpu.py -x -q > gpurun_out/c2_pytest.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/c2_pytest.log
timeout 200 python tools/trace_step.py > gpurun_out/c2_trace_n1.log 2>&1; grep TRACE gpurun_out/c2_trace_n1.log
timeout 300 $TR --master-port 29621 tools/trace_step.py > gpurun_out/c2_trace_n2.log 2>&1; grep TRACE gpurun_out/c2_trace_n2.log
TFDL_AR_ALGO=oneshot timeout 300 $TR --master-port 29622 tools/trace_step.py > gpurun_out/c2_trace_n2_oneshot.log 2>&1; grep TRACE gpurun_out/c2_trace_n2_oneshot.log
# two independent single-GPU runs side by side: does a second busy process alone slow a rank down?
(CUDA_VISIBLE_DEVICES=0 timeout 200 python bench.py --steps 400 --warmup 20 > gpurun_out/c2_solo0.json 2>/dev/null &
 CUDA_VISIBLE_DEVICES=1 timeout 200 python bench.py --steps 400 --warmup 20 > gpurun_out/c2_solo1.json 2>/dev/null & wait)
cut -c1-200 gpurun_out/c2_solo0.json gpurun_out/c2_solo1.json
nproc; cat /proc/cpuinfo | grep "model name" | head -1
