#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
export TFDL_TRACE_MASK=0x41
timeout 200 python tools/trace_step.py 2>&1 | grep TRACE
timeout 300 $TR --master-port 29621 tools/trace_step.py 2>&1 | grep TRACE
TFDL_PDL=0 timeout 300 $TR --master-port 29623 tools/trace_step.py 2>&1 | grep TRACE
TFDL_AR_OVERLAP=0 timeout 300 $TR --master-port 29624 tools/trace_step.py 2>&1 | grep TRACE
