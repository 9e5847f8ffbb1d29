#!/bin/bash
# round 2, first GPU pass: full-model parity, DCN vs the reference extension, comparators, then the whole GPU suite
set -u
mkdir -p gpurun_out
rm -f gpurun_out/fullmodel_parity.jsonl
timeout 1500 python -m pytest tests/test_gpu_fullmodel.py tests/test_gpu_dcn_reference.py tests/test_gpu_volumes.py -m gpu -q -s -x > gpurun_out/r2_new_tests.log 2>&1; echo "new tests rc=$?"
tail -25 gpurun_out/r2_new_tests.log
timeout 600 python tools/time_dcn_ref.py > gpurun_out/r2_dcn_ref.log 2>&1; echo "dcn_ref rc=$?"; cat gpurun_out/r2_dcn_ref.log | tail -8
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
