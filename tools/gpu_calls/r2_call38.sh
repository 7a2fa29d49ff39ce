#!/bin/bash
# 2 GPUs: the driver's bench form (K = 20, W = 5) with the prefetched gather, and the TF_CONFIG two-worker fit test
mkdir -p gpurun_out
TFDL_FIT_PROFILE=1 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 20 --warmup 5 2>gpurun_out/c38_n2.err > gpurun_out/c38_n2.json; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/c38_n2.json").read().strip().splitlines()[-1])
print("N=2 device us/step", d["ms_per_step"]*1e3, "e2e us/step", d["e2e"]["ms_per_step"]*1e3, "e2e img/s", d["e2e"]["value"], "loss", d["e2e"]["final_loss"])
PY
grep "fit profile" gpurun_out/c38_n2.err | tail -2
timeout 200 python -m pytest tests/test_multigpu.py -m gpu -x -q -k "tf_config_workers" > gpurun_out/c38_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/c38_pytest.log | cut -c1-300
