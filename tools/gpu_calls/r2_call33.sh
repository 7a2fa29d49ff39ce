#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
for m in resnet50 bert; do
timeout 400 $TR --master-port 29791 bench.py --model $m --gpus 2 --steps 10 --warmup 5 2>/dev/null > gpurun_out/c33_${m}_n2.json
python -c "import json; d=json.loads(open('gpurun_out/c33_${m}_n2.json').read().strip().splitlines()[-1]); print('$m', d['n_gpus'], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['gpu_launches'], len(d['config']['buckets'] or []))"
done
timeout 300 python bench.py --model resnet50 --steps 10 --warmup 5 2>/dev/null > gpurun_out/c33_resnet50_n1.json
python -c "import json; d=json.loads(open('gpurun_out/c33_resnet50_n1.json').read().strip().splitlines()[-1]); print('resnet50', d['n_gpus'], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['gpu_launches'])"
