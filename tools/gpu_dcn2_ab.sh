#!/bin/bash
# A/B of the channels-last DCN forward kernel (tuning build: env hooks SF_DCN2_NA / SF_DCN2_NST / SF_DCN2_COMPACT)
export SF_TUNING_LIB=1
run() { echo "== $*"; env "$@" timeout 100 python tools/time_dcn2.py ${SH:-dcn_128} 2>&1 | tail -3 | cut -c1-200; }
SH="dcn_128 dcn_128x8 dcn_256" run A=base
SH="dcn_128 dcn_128x8" run SF_DCN2_COMPACT=1
SH="dcn_128 dcn_128x8" run SF_DCN2_COMPACT=1 SF_DCN2_NA=3
SH="dcn_128 dcn_128x8" run SF_DCN2_COMPACT=1 SF_DCN2_NA=4
SH="dcn_128 dcn_128x8 dcn_256" run A=base
