#!/bin/bash
# timing of the channels-last DCN forward kernel (tuning build: env hooks SF_DCN2_NA / SF_DCN2_NST / SF_DCN2_COMPACT)
export SF_TUNING_LIB=1
run() { echo "== $*"; env "$@" timeout 100 python tools/time_dcn2.py ${SH:-dcn_128} 2>&1 | tail -5 | cut -c1-200; }
SH="dcn_32 dcn_128 dcn_128x8 dcn_256" run A=base
