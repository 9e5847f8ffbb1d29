#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_precision.py tests/test_gpu_gather.py -m gpu -q -s -x 2>&1 | grep -v Warning | tail -15
timeout 900 python tools/exp_precision.py 2>/dev/null | tee gpurun_out/exp_precision.jsonl
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
