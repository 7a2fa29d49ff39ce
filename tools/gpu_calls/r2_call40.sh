#!/bin/bash
# op-determinism switch on the GPU: two seeded runs bitwise equal (tests/test_determinism.py), plus the two files whose
# code paths the switch touches
mkdir -p gpurun_out
timeout 280 python -m pytest tests/test_determinism.py tests/test_native_engine_gpu.py -m gpu -x -q > gpurun_out/c40_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/c40_pytest.log | cut -c1-400
