#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
export TFDL_TRACE_MASK=0x41
for e in 0 1 2; do
TFDL_EXP=$e timeout 200 python tools/trace_step.py 2>&1 | grep TRACE
TFDL_EXP=$e timeout 300 $TR --master-port 2962$e tools/trace_step.py 2>&1 | grep TRACE
done
