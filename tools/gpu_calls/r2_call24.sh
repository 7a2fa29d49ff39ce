#!/bin/bash
export TFDL_FIT_PROFILE=1
timeout 300 python bench.py --steps 20 --warmup 5 2>&1 >/dev/null | grep "fit profile"
timeout 300 python bench.py --steps 20 --warmup 5 2>&1 >/dev/null | grep "fit profile"
