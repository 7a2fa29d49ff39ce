#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu.py tests/test_zz_no_alloc.py tests/test_conv_gpu.py -x -q > gpurun_out/c16_pytest.log 2>&1; echo "pytest rc=$?"; tail -30 gpurun_out/c16_pytest.log | cut -c1-250
timeout 300 python bench.py --model resnet50 --steps 10 --warmup 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('resnet50', d['value'], d['ms_per_step'])"
