#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/c36_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/c36_pytest.log | cut -c1-200
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py 2>gpurun_out/c36_bench.err > gpurun_out/c36_bench.json; echo "bench rc=$?"; cut -c1-1500 gpurun_out/c36_bench.json
python bench.py --impl reference
