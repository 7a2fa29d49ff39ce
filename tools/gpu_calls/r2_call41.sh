#!/bin/bash
# last call of the round: the GPU suite on the final tree (the determinism test, three fresh processes, ran in call 40)
mkdir -p gpurun_out
timeout 30 python -m pytest tests -m gpu -x -q --deselect tests/test_determinism.py > gpurun_out/c41_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/c41_pytest.log | cut -c1-300
