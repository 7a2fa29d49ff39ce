#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/conv_bench.py --quick > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:conv_igemm -s 6 -c 1 -o gpurun_out/ncu_conv_igemm2 python tools/conv_bench.py --quick > gpurun_out/ncu_conv.log 2>&1; echo "ncu rc=$?"
