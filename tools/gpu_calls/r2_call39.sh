#!/bin/bash
# final single-GPU validation: host topology, GPU suite, smoke, the driver's bench form, reference arm
mkdir -p gpurun_out
python tools/cpu_topology.py 2>&1 | tail -1 | cut -c1-1800 | tee gpurun_out/c39_topology.json
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/c39_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/c39_pytest.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
TFDL_FIT_PROFILE=1 timeout 200 python bench.py --gpus 1 --steps 20 --warmup 5 2>gpurun_out/c39_bench.err > gpurun_out/c39_bench.json; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/c39_bench.json").read().strip().splitlines()[-1])
print("device us/step", d["ms_per_step"]*1e3, "e2e us/step", d["e2e"]["ms_per_step"]*1e3, "e2e img/s", d["e2e"]["value"])
PY
grep "fit profile" gpurun_out/c39_bench.err | tail -1
python bench.py --impl reference
