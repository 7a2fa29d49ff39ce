#!/bin/bash
# final 8-GPU run of the round: multi-GPU tests, scaling table (driver-style short run and long run), small-size sweep
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_multigpu.py -x -q > gpurun_out/c28_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/c28_pytest.log | cut -c1-200
for n in 1 2 4 8; do
  for st in 400 20; do
    if [ $n = 1 ]; then timeout 300 python bench.py --steps $st --warmup 10 2>gpurun_out/c28_n${n}_$st.err > gpurun_out/c28_n${n}_$st.json
    else TFDL_FIT_PROFILE=1 timeout 300 $TR --nproc-per-node $n --master-port 2975$n bench.py --gpus $n --steps $st --warmup 10 2>gpurun_out/c28_n${n}_$st.err > gpurun_out/c28_n${n}_$st.json; fi
    python -c "import json; d=json.loads(open('gpurun_out/c28_n${n}_$st.json').read().strip().splitlines()[-1]); print(d['n_gpus'], d['steps'], d['value'], d['ms_per_step'], d['config']['allreduce'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], d['clocks'])"
  done
done
grep "fit profile" gpurun_out/c28_n8_400.err | tail -2 | cut -c1-250
timeout 600 $TR --nproc-per-node 8 --master-port 29761 tools/allreduce_sweep.py --max-mb 8 --no-sum --tag _small > gpurun_out/c28_sweep_n8.log 2>&1; grep '^{' gpurun_out/c28_sweep_n8.log | python -c "
import sys,json
for l in sys.stdin:
    r=json.loads(l); print(r['bytes'], ' '.join(f'{k[6:-3]}={r[k]:.1f}' for k in r if k.startswith('fused_') and k.endswith('_us')), 'nccl=%.1f'%r['nccl_us'])"
