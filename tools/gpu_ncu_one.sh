#!/bin/bash
# usage: gpu_ncu_one.sh <kernel regex> <out name> <python script + args...>
set -u
K=$1; O=$2; shift 2
mkdir -p gpurun_out
timeout 300 python "$@" > gpurun_out/$O.plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/$O.plain.log; exit 1; }
timeout 240 ncu --set full --clock-control none --import-source on -k regex:$K -c 1 -f -o gpurun_out/$O python "$@" > gpurun_out/$O.ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/$O.ncu.log; ls -la gpurun_out/$O.ncu-rep
