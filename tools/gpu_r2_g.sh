#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; grep -v Warning gpurun_out/pytest_gpu.log | tail -12
timeout 600 python -m pytest tests/test_gpu_volumes.py tests/test_gpu_fullmodel.py tests/test_gpu_precision.py -m gpu -q -s 2>&1 | grep "chain:\|variance window\|zero-offset\|cost volume C\|'variant'" | cut -c1-600
