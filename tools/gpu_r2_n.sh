#!/bin/bash
# multi-GPU pass: N = $1 ; $2 = "full" also runs the all-rank and NCCL gather variants
set -u
N=${1:-2}; MODE=${2:-short}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 300 $TR tools/check_gather.py > gpurun_out/check_gather_n$N.log 2>&1; echo "check_gather rc=$?"; grep "gather to" gpurun_out/check_gather_n$N.log
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n${N}_s20.json 2> gpurun_out/bench_n$N.err; echo "bench s20 rc=$?"; tail -2 gpurun_out/bench_n$N.err
timeout 600 $TR bench.py --gpus $N --no-configs > gpurun_out/bench_n${N}.json 2>> gpurun_out/bench_n$N.err; echo "bench rc=$?"
if [ "$MODE" = "full" ]; then
timeout 600 $TR bench.py --gpus $N --no-configs --gather nccl > gpurun_out/bench_n${N}_nccl.json 2>> gpurun_out/bench_n$N.err; echo "bench nccl rc=$?"
timeout 600 $TR bench.py --gpus $N --no-configs --gather all > gpurun_out/bench_n${N}_all.json 2>> gpurun_out/bench_n$N.err; echo "bench all rc=$?"
fi
for f in gpurun_out/bench_n${N}_s20.json gpurun_out/bench_n${N}.json; do [ -f $f ] && grep '^{' $f | python -c "
import json,sys
j=json.loads(sys.stdin.read()); r=j['roofline']; c=j.get('configs',{}).get('c5_bands') or {}
print('$f', 'value',round(j['value']), 'us/step', round(j['ms_per_step']*1e3,2), 'cv us', round(r['us_per_launch_before_timed_region'],2), round(r['us_per_launch_after_timed_region'],2), 'serial', round(r['serial_step_us'],2), 'blocks', j['timed_blocks'], 'e2e', round(j['e2e']['value']), j['clocks']['reasons'], 'checked', j.get('gather_steps_checked_in_warmup'), 'c5_bands us', c.get('us'), c.get('upsampled_bit_exact_vs_whole_image'), c.get('error'))
"; done
