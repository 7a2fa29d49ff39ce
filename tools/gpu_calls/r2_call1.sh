#!/bin/bash
# round-2 GPU call 1 (2 GPUs): push all-reduce correctness + flagship step with/without it + small sweep
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_multigpu.py tests/test_native_engine_gpu.py -x -q > gpurun_out/c1_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/c1_pytest.log
timeout 300 python bench.py --steps 400 --warmup 20 > gpurun_out/c1_bench_n1.json 2> gpurun_out/c1_bench_n1.err; tail -1 gpurun_out/c1_bench_n1.json
for cfg in "default" "oneshot" "push_nooverlap" "push2dense"; do
  case $cfg in
    default) E="";;
    oneshot) E="TFDL_AR_ALGO=oneshot";;
    push_nooverlap) E="TFDL_AR_OVERLAP=0";;
    push2dense) E="TFDL_AR_PUSH_MAX=400000";;
  esac
  env $E timeout 300 $TR --master-port 29611 bench.py --gpus 2 --steps 400 --warmup 20 > gpurun_out/c1_bench_n2_$cfg.json 2> gpurun_out/c1_bench_n2_$cfg.err
  tail -1 gpurun_out/c1_bench_n2_$cfg.json | cut -c1-400
done
timeout 600 $TR --master-port 29612 tools/allreduce_sweep.py --max-mb 8 --no-sum --tag _c1 > gpurun_out/c1_sweep_n2.log 2>&1
tail -20 gpurun_out/c1_sweep_n2.log
