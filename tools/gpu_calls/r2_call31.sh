#!/bin/bash
timeout 900 python -m pytest tests/test_models_gpu.py tests/test_conv_gpu.py tests/test_native_engine_gpu.py -x -q 2>&1 | tail -12 | cut -c1-250
