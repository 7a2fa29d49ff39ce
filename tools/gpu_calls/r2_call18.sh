#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/conv_bench.py > gpurun_out/conv_bench.md 2> gpurun_out/conv_bench.err; echo "conv_bench rc=$?"; cat gpurun_out/conv_bench.md; tail -3 gpurun_out/conv_bench.err
timeout 300 python tools/conv_bench.py --quick > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:conv_igemm -s 6 -c 3 -o gpurun_out/ncu_conv_igemm python tools/conv_bench.py --quick > gpurun_out/ncu_conv.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_conv.log
