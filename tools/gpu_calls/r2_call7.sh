#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_native_engine_gpu.py tests/test_multigpu.py -x -q > gpurun_out/c7_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/c7_pytest.log
for st in 20 400; do
timeout 300 python bench.py --steps $st --warmup 5 2>gpurun_out/c7_n1_$st.err | tee gpurun_out/c7_n1_$st.json | cut -c1-200
timeout 300 $TR --master-port 29641 bench.py --gpus 2 --steps $st --warmup 5 2>gpurun_out/c7_n2_$st.err | tee gpurun_out/c7_n2_$st.json | cut -c1-200
done
timeout 300 python bench.py --impl standin --steps 400 --warmup 5 2>gpurun_out/c7_s1.err | tee gpurun_out/c7_s1.json | cut -c1-200
timeout 300 $TR --master-port 29642 bench.py --impl standin --gpus 2 --steps 400 --warmup 5 2>gpurun_out/c7_s2.err | tee gpurun_out/c7_s2.json | cut -c1-200
tail -5 gpurun_out/c7_*.err | cut -c1-300
