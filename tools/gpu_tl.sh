#!/bin/bash
python -m pytest tests/test_gpu_volumes.py tests/test_gpu_bands.py -m gpu -q -x 2>&1 | tail -2
python tools/timeline_cv.py 2>&1 | tail -9 | cut -c1-700
for sh in "8 256 68 120 24" "1 256 248 360 48" "1 256 32 64 24" "4 256 136 240 48" "8 192 68 120 24" "8 320 68 120 24"; do timeout 120 python tools/time_cv.py $sh; done
python bench.py --steps 20 --warmup 3 --no-configs | python -c "
import json,sys
j=json.loads(sys.stdin.read()); r=j['roofline']; print('bench K=20: value',round(j['value']),'us/step',round(j['ms_per_step']*1e3,2),'cv us',round(r['us_per_launch'],2),'frac',round(r['frac'],3),'serial',round(r['serial_step_us'],2), 'eager', r['eager_api_step_us'], j['clocks']['reasons'])"
python bench.py --no-configs | python -c "
import json,sys
j=json.loads(sys.stdin.read()); r=j['roofline']; print('bench K=300: value',round(j['value']),'us/step',round(j['ms_per_step']*1e3,2),'cv us',round(r['us_per_launch'],2),'frac',round(r['frac'],3),'serial',round(r['serial_step_us'],2), j['clocks'])"
