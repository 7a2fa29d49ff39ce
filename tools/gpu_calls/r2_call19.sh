#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_conv_gpu.py -x -q 2>&1 | tail -3
timeout 600 python tools/conv_bench.py > gpurun_out/conv_bench.md 2> gpurun_out/conv_bench.err; cat gpurun_out/conv_bench.md
timeout 300 python bench.py --model resnet50 --steps 10 --warmup 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('resnet50', d['value'], d['ms_per_step'])"
