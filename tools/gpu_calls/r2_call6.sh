#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_native_engine_gpu.py tests/test_multigpu.py -x -q > gpurun_out/c6_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/c6_pytest.log
for k in 8 1; do
TFDL_STEPS_PER_GRAPH=$k timeout 300 python bench.py --steps 400 --warmup 20 2>gpurun_out/c6_n1_$k.err | tee gpurun_out/c6_n1_$k.json | cut -c1-260
TFDL_STEPS_PER_GRAPH=$k timeout 300 $TR --master-port 2963$k bench.py --gpus 2 --steps 400 --warmup 20 2>gpurun_out/c6_n2_$k.err | tee gpurun_out/c6_n2_$k.json | cut -c1-260
done
timeout 300 python bench.py --steps 20 --warmup 3 2>/dev/null | tee gpurun_out/c6_n1_20steps.json | cut -c1-260
