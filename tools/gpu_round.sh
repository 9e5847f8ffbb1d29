#!/bin/bash
# One GPU-box pass: parity tests, bench, kernel timings, ncu launch list + full capture of the top kernel.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2>&1; cat gpurun_out/bench_ref.json
python tools/time_cv.py > gpurun_out/time_cv.log 2>&1; cat gpurun_out/time_cv.log
python tools/time_attn.py > gpurun_out/time_attn.log 2>&1; cat gpurun_out/time_attn.log
python tools/time_corr1d.py > gpurun_out/time_corr1d.log 2>&1; cat gpurun_out/time_corr1d.log
python tools/time_ops.py > gpurun_out/time_ops.log 2>&1; tail -20 gpurun_out/time_ops.log
python tools/time_refiner.py 8 > gpurun_out/time_refiner.log 2>&1; tail -3 gpurun_out/time_refiner.log
CMD="python bench.py --steps 20 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:corr_volume -s 5 -c 2 -f -o gpurun_out/prof_cv $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out
