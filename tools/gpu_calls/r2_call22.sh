#!/bin/bash
# 8 GPUs: dense-bucket CTA count variants, scaling table, step trace, single-process launch, NVLink reference copy
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for blocks in 64 148 32; do
  TFDL_DENSE_AR_BLOCKS=$blocks timeout 300 $TR --nproc-per-node 8 --master-port 2971${blocks:0:1} bench.py --gpus 8 --steps 400 --warmup 10 2>/dev/null > gpurun_out/c22_n8_b$blocks.json
  python -c "import json; d=json.loads(open('gpurun_out/c22_n8_b$blocks.json').read().strip().splitlines()[-1]); print('blocks $blocks', d['n_gpus'], d['value'], d['ms_per_step'], d['config']['allreduce'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'])"
done
TFDL_DENSE_AR_BLOCKS=148 timeout 300 $TR --nproc-per-node 8 --master-port 29721 tools/trace_step.py 2>&1 | grep TRACE | head -1 | cut -c1-500
timeout 300 python tools/bench_single_process.py --steps 400 2>&1 | grep -E "SINGLE_PROCESS|Error" | cut -c1-300
timeout 120 python tools/nvlink_peaks.py 2>&1 | grep NVLINK
for n in 1 2 4; do
  if [ $n = 1 ]; then timeout 300 python bench.py --steps 400 --warmup 10 2>/dev/null > gpurun_out/c22_n$n.json
  else timeout 300 $TR --nproc-per-node $n --master-port 2973$n bench.py --gpus $n --steps 400 --warmup 10 2>/dev/null > gpurun_out/c22_n$n.json; fi
  python -c "import json; d=json.loads(open('gpurun_out/c22_n$n.json').read().strip().splitlines()[-1]); print(d['n_gpus'], d['value'], d['ms_per_step'], d['config']['allreduce'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'])"
done
