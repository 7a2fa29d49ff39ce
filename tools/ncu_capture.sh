#!/bin/bash
# One `ncu --set full` capture of every kernel of the flagship training step (eager launches so the kernels are
# visible by name), only after the same command has exited 0 without ncu. One GPU. Output: gpurun_out/step_full.ncu-rep
# Read it on the build box with tools/ncu_summarise.sh.
set -u
export TFDL_USE_GRAPH=0
mkdir -p gpurun_out
python bench.py --steps 5 --warmup 3 > gpurun_out/plain_eager.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_eager.log; exit 1; }
ncu --set full --clock-control none --import-source on \
    -k regex:'conv12|dense_block|conv21|conv1_wgrad|fused_ar' -s 20 -c 5 \
    -f -o gpurun_out/step_full python bench.py --steps 5 --warmup 3 > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/step_full.ncu-rep
