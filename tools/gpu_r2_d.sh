#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_dcn_reference.py tests/test_gpu_ops.py tests/test_gpu_precision.py tests/test_gpu_bands.py -m gpu -q -x -k "dcn or precision or band" 2>&1 | grep -v Warning | tail -8
P='import tools.time_dcn_ref as t; r=t.run(shapes=t.SHAPES[:1], quiet=True)["dcn_16"]; print({k: round(v,2) for k,v in r.items() if k.endswith("_us") or k.startswith("max")})'
echo "product (quad, PPT=1)"; python -c "$P"
echo "tuning pair PPT=2"; SF_TUNING_LIB=1 SF_DCN_PAIR=2 python -c "$P"
