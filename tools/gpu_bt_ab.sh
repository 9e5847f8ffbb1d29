#!/bin/bash
# A/B of the cost-volume backward plan in the tuning build (env hooks SF_BT_*): bench grid
export SF_TUNING_LIB=1
run() { echo "== $*"; env "$@" timeout 60 python tools/time_cv_bwd.py tc ${DD:-24} 2>&1 | grep -E "^\{|Error" | cut -c1-150; }
DD=24,33,48,65 run A=base
DD=24,33,48,65 run SF_BT_NO=2
