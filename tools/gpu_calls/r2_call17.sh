#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu.py -x -q -k "single_process or killed" > gpurun_out/c17_pytest.log 2>&1; echo "pytest rc=$?"; tail -40 gpurun_out/c17_pytest.log | cut -c1-300
timeout 300 python tools/bench_single_process.py --steps 400 2>&1 | grep -E "SINGLE_PROCESS|Error|error" | cut -c1-300
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus 2 --steps 400 --warmup 10 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('torchrun', d['n_gpus'], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'])"
