#!/bin/bash
timeout 600 python -m pytest tests/test_native_engine_gpu.py -x -q 2>&1 | tail -2
for st in 20 400; do
timeout 300 python bench.py --steps $st --warmup 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['n_gpus'], d['steps'], d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['e2e']['final_loss'])"
done
