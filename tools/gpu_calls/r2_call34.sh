#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu.py -x -q -k "all_visible" > gpurun_out/c34_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/c34_pytest.log | cut -c1-250
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29801 tools/allreduce_sweep.py --min-kb 65536 --max-mb 256 --algos twoshot,multimem --tag _large > gpurun_out/c34_sweep.log 2>&1; grep '^{' gpurun_out/c34_sweep.log | python -c "
import sys,json
for l in sys.stdin:
    r=json.loads(l); print(r['bytes'], ' '.join(f'{k[:-6]}={r[k]:.0f}' for k in r if k.endswith('_busbw')))"
