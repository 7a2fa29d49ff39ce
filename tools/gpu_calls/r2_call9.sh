#!/bin/bash
# 8-GPU call: multi-GPU tests, flagship bench at 1/2/4/8, step trace at 8, all-reduce sweeps at 8 and 4
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_multigpu.py -x -q > gpurun_out/c9_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/c9_pytest.log
for n in 1 2 4 8; do
  if [ $n = 1 ]; then timeout 300 python bench.py --steps 400 --warmup 10 2>gpurun_out/c9_n$n.err > gpurun_out/c9_n$n.json
  else timeout 300 $TR --nproc-per-node $n --master-port 2965$n bench.py --gpus $n --steps 400 --warmup 10 2>gpurun_out/c9_n$n.err > gpurun_out/c9_n$n.json; fi
  python -c "import json,sys; d=json.loads(open('gpurun_out/c9_n$n.json').read().strip().splitlines()[-1]); print(d['n_gpus'], d['value'], d['ms_per_step'], d['config']['allreduce'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'])"
done
TFDL_TRACE_MASK=0x41 timeout 300 $TR --nproc-per-node 8 --master-port 29661 tools/trace_step.py 2>&1 | grep TRACE | cut -c1-300
timeout 300 $TR --nproc-per-node 8 --master-port 29662 tools/trace_step.py 2>&1 | grep TRACE | head -2 | cut -c1-500
timeout 900 $TR --nproc-per-node 8 --master-port 29663 tools/allreduce_sweep.py --max-mb 1024 > gpurun_out/c9_sweep_n8.log 2>&1; tail -22 gpurun_out/c9_sweep_n8.log
timeout 600 $TR --nproc-per-node 4 --master-port 29664 tools/allreduce_sweep.py --max-mb 256 > gpurun_out/c9_sweep_n4.log 2>&1; tail -20 gpurun_out/c9_sweep_n4.log
