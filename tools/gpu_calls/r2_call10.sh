#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_multigpu.py -x -q > gpurun_out/c10_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/c10_pytest.log
timeout 300 $TR --master-port 29671 bench.py --gpus 2 --steps 400 --warmup 10 2>gpurun_out/c10_n2.err > gpurun_out/c10_n2.json
python -c "import json,sys; d=json.loads(open('gpurun_out/c10_n2.json').read().strip().splitlines()[-1]); print(d['n_gpus'], d['value'], d['ms_per_step'], d['config']['allreduce'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'])"
timeout 600 $TR --master-port 29672 tools/allreduce_sweep.py --max-mb 8 --no-sum --algos push,push2,oneshot --tag _c10 > gpurun_out/c10_sweep_n2.log 2>&1; grep '^{' gpurun_out/c10_sweep_n2.log | python -c "
import sys,json
for l in sys.stdin:
    r=json.loads(l); print(r['bytes'], ' '.join(f'{k[6:-3]}={r[k]:.1f}' for k in r if k.startswith('fused_') and k.endswith('_us')), 'nccl=%.1f'%r['nccl_us'])"
