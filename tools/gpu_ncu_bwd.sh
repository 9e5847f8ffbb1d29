#!/bin/bash
set -u
mkdir -p gpurun_out
bash tools/gpu_ncu_one.sh dcn_bwd_gcol_tc dcn_gcol tools/run_dcn_bwd_once.py
bash tools/gpu_ncu_one.sh dcn_bwd_scatter dcn_scatter tools/run_dcn_bwd_once.py
bash tools/gpu_ncu_one.sh dcn_bwd_weight_tc dcn_wgrad tools/run_dcn_bwd_once.py
