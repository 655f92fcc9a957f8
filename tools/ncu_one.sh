#!/bin/bash
# One ncu --set full capture of a kernel of tools/profile_target.py, raw CSV into gpurun_out/.
#   ncu_one.sh <tag> <kernel regex> <skip> <count> <profile_target args...>
set -u
tag=$1; K=$2; skip=$3; cnt=$4; shift 4
python tools/profile_target.py "$@" > gpurun_out/one_${tag}.plain.log 2>&1 || { echo "plain run failed"; cat gpurun_out/one_${tag}.plain.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:$K" -s $skip -c $cnt -f \
    -o gpurun_out/one_${tag} python tools/profile_target.py "$@" > gpurun_out/one_${tag}.ncu.log 2>&1
ncu -i gpurun_out/one_${tag}.ncu-rep --page raw --csv > gpurun_out/one_${tag}_raw.csv 2>/dev/null
ncu -i gpurun_out/one_${tag}.ncu-rep --page details --csv > gpurun_out/one_${tag}_details.csv 2>/dev/null
rm -f gpurun_out/one_${tag}.ncu-rep
tail -2 gpurun_out/one_${tag}.plain.log
