#!/bin/bash
# Per-kernel device times of the training step (eager program so kernel names are visible).
# Usage under gpurun: bash tools/launch_times.sh  -> gpurun_out/launches.csv
export TFDL_USE_GRAPH=0
python bench.py --steps 20 --warmup 5 > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 150 -c 20 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 20 --warmup 5 > gpurun_out/ncu.log 2>&1
tail -1 gpurun_out/ncu.log | cut -c1-200
