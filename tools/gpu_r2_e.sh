#!/bin/bash
cd $GRAFT_REPO_ROOT
echo "--- PPT=2"; timeout 300 compute-sanitizer --tool memcheck --print-limit 5 python tools/dcn_one.py 1 16 16 34 60 2>&1 | grep -v "^=========     at\|^=========     by\|Host Frame\|libtorch\|python3\|libc.so\|libpython" | head -40
echo "--- PPT=1"; SF_TUNING_LIB=1 SF_DCN_PPT=1 timeout 300 python tools/dcn_one.py 1 16 16 34 60 2>&1 | tail -3
