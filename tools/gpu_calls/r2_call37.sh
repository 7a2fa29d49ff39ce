#!/bin/bash
# prefetched gather in fit(): GPU suite, then the driver's bench form (K = 20, W = 5) with and without it
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/c37_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/c37_pytest.log | cut -c1-300
for pf in 1 0; do
  TFDL_FIT_PROFILE=1 TFDL_FIT_PREFETCH=$pf timeout 200 python bench.py --gpus 1 --steps 20 --warmup 5 2>gpurun_out/c37_pf$pf.err > gpurun_out/c37_pf$pf.json; echo "bench pf=$pf rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/c37_pf$pf.json").read().strip().splitlines()[-1])
print("pf=$pf device us/step", d["ms_per_step"]*1e3, "e2e us/step", d["e2e"]["ms_per_step"]*1e3, "e2e img/s", d["e2e"]["value"])
PY
  grep "fit profile" gpurun_out/c37_pf$pf.err | tail -1
done
