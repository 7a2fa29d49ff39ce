#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_models_gpu.py tests/test_conv_gpu.py -x -q > gpurun_out/c12_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/c12_pytest.log | cut -c1-220
for m in resnet50 bert; do
timeout 600 python bench.py --model $m --steps 10 --warmup 5 2>gpurun_out/c12_${m}_n1.err > gpurun_out/c12_${m}_n1.json; tail -3 gpurun_out/c12_${m}_n1.err | cut -c1-300
python -c "import json; d=json.loads(open('gpurun_out/c12_${m}_n1.json').read().strip().splitlines()[-1]); print('$m', d['n_gpus'], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['gpu_launches'], d['config']['buckets'])"
timeout 600 $TR --master-port 29681 bench.py --model $m --gpus 2 --steps 10 --warmup 5 2>gpurun_out/c12_${m}_n2.err > gpurun_out/c12_${m}_n2.json; tail -3 gpurun_out/c12_${m}_n2.err | cut -c1-300
python -c "import json; d=json.loads(open('gpurun_out/c12_${m}_n2.json').read().strip().splitlines()[-1]); print('$m', d['n_gpus'], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['gpu_launches'], d['config']['buckets'])"
done
TFDL_CONV=torch timeout 600 python bench.py --model resnet50 --steps 10 --warmup 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('resnet50 cudnn-conv', d['value'], d['ms_per_step'])"
