#!/bin/bash
# ncu --set full captures of the round-2 kernels at their headline shapes (one launch each), summaries by tools/ncu_summary.py
set -u
mkdir -p gpurun_out
bash tools/gpu_ncu_one.sh corr_volume_bwd_tc cv_bwd tools/time_cv_bwd.py once 0
bash tools/gpu_ncu_one.sh dcn_tc2_kernel dcn_tc2 tools/run_dcn_once.py
bash tools/gpu_ncu_one.sh epipolar_attn_tc attn tools/time_attn.py 8 128 96 312 48
