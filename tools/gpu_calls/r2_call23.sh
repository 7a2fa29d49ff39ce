#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_native_engine_gpu.py tests/test_multigpu.py -x -q > gpurun_out/c23_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/c23_pytest.log | cut -c1-250
for st in 20 400; do
timeout 300 python bench.py --steps $st --warmup 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['n_gpus'], d['steps'], d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['e2e']['final_loss'])"
timeout 300 $TR --master-port 29741 bench.py --gpus 2 --steps $st --warmup 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['n_gpus'], d['steps'], d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['e2e']['final_loss'])"
done
