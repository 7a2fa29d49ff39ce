#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/c32_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/c32_pytest.log | cut -c1-250
timeout 300 python bench.py --steps 20 --warmup 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['n_gpus'], d['steps'], d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['gpu_launches'])"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29781 bench.py --gpus 2 --steps 20 --warmup 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['n_gpus'], d['steps'], d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'])"
