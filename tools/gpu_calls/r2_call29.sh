#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_multigpu.py -x -q -k single_process > gpurun_out/c29_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/c29_pytest.log | cut -c1-200
for st in 20 20 400; do
timeout 300 $TR --nproc-per-node 4 --master-port 29771 bench.py --gpus 4 --steps $st --warmup 10 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['n_gpus'], d['steps'], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'])"
done
