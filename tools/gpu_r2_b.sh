#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gather.py tests/test_gpu_fullmodel.py tests/test_gpu_volumes.py -m gpu -q -x > gpurun_out/r2_b_tests.log 2>&1; echo "tests rc=$?"; tail -15 gpurun_out/r2_b_tests.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -5 gpurun_out/bench.err; python -c "
import json; j=json.load(open('gpurun_out/bench.json')); c=j.pop('configs'); print(json.dumps(j)); 
for k,v in c.items(): print(k, json.dumps(v))"
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2>gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json; tail -3 gpurun_out/bench_ref.err
timeout 300 python bench.py --steps 20 --warmup 3 --no-configs > gpurun_out/bench_s20.json 2>> gpurun_out/bench.err; cat gpurun_out/bench_s20.json
