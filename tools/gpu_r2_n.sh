#!/bin/bash
# multi-GPU pass: N = $1
set -u
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 300 python -m pytest tests/test_gpu_gather.py -m gpu -q -x 2>&1 | tail -3
timeout 300 $TR tools/check_gather.py > gpurun_out/check_gather_n$N.log 2>&1; echo "check_gather rc=$?"; tail -4 gpurun_out/check_gather_n$N.log
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n${N}_s20.json 2> gpurun_out/bench_n$N.err; echo "bench s20 rc=$?"; tail -3 gpurun_out/bench_n$N.err; cat gpurun_out/bench_n${N}_s20.json
timeout 600 $TR bench.py --gpus $N --no-configs > gpurun_out/bench_n${N}.json 2>> gpurun_out/bench_n$N.err; echo "bench rc=$?"; cat gpurun_out/bench_n${N}.json
timeout 600 $TR bench.py --gpus $N --no-configs --gather all > gpurun_out/bench_n${N}_all.json 2>> gpurun_out/bench_n$N.err; echo "bench all rc=$?"; cat gpurun_out/bench_n${N}_all.json
timeout 600 $TR bench.py --gpus $N --no-configs --gather nccl > gpurun_out/bench_n${N}_nccl.json 2>> gpurun_out/bench_n$N.err; echo "bench nccl rc=$?"; cat gpurun_out/bench_n${N}_nccl.json
