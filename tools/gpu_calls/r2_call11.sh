#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_conv_gpu.py -x -q > gpurun_out/c11_conv.log 2>&1; echo "conv pytest rc=$?"; tail -30 gpurun_out/c11_conv.log | cut -c1-250
bash tools/gpu_calls/r2_call10.sh
